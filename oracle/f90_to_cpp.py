#!/usr/bin/env python
"""Syntactic translator of the Fortran 90 subset the reference's integrator is written in (fortran/raytracer.f95,
util.f95, bmodel_dipole.f95, constants.f95) to C++ over oracle/ref_stubs/f90rt.h.

TEST INFRASTRUCTURE (oracle/): no Fortran compiler exists here, so the reference's own source is translated where it lies
(/root/reference, read at build time only) and compiled into oracle/_ref/ -- the translated text never enters the
repository.  Unlike the FORTRAN 77 files (f77_to_cpp.py), this source uses array expressions: the translation is purely
syntactic (declarations -> Vec / Mat objects, sections -> views, statements one to one) and every array semantic lives in
f90rt.h, which states the choices a compiler makes (element order of SUM / DOT_PRODUCT, MAX with a NaN, CMPLX without
KIND).  Dummy procedures (funcPlasmaParams, funcStopConditions, their data argument) are bound to hooks of the test
harness; `csvd` (LAPACK zgesvd in the reference) is a hook as well.
Usage: f90_to_cpp.py out.cpp file.f95 [file.f95 ...]
"""
import os
import re
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from f77_to_cpp import read_statements_free, split_top, top_level_eq  # noqa: E402

HOOK_SUBS = {"SM_TO_GSM_D": "RT_sm_to_gsm", "GSM_TO_SM_D": "RT_gsm_to_sm", "TSY_RECALC": "RT_tsy_recalc", "IGRF_GSM": "RT_igrf_gsm",
             "T04_S": "RT_t04_s", "DENS": "RT_dens"}
DUMMY_PROCS = {"FUNCPLASMAPARAMS", "FUNCPLASMAPARAMSDATA", "FUNCSTOPCONDITIONS", "DAT"}  # DAT: the state argument of setup (its components are module variables here)
KEEP_UNITS = {"FUNCPLASMAPARAMS", "DISPERSION_RELATION", "STIX_PARAMETERS", "DISPERSION_RELATION_DFDK", "DISPERSION_RELATION_DFDW", "DISPERSION_RELATION_DFDX",
              "RAYTRACER_EVALRHS", "RAYTRACER_STOPCONDITIONS", "FORM_DISPERSION_MATRIX", "IS_RIGHT_HANDED", "SOLVE_DISPERSION_RELATION", "RK4",
              "RK45", "RAYTRACER_RUN", "IJK2N", "POINT2XYZ", "TRICUBIC_GET_COEFF", "TRICUBIC_EVAL", "TRICUBIC_GET_COEFF_STACKED", "TRICUBIC_INTERPOLATE_AT", "TRICUBIC_COMPUTE_FINITE_DIFFERENCE_DERIVATIVES", "CARTESIAN_TO_SPHERICAL", "SPHERICAL_TO_CARTESIAN_VEC", "CARTESIAN_TO_SPHERICAL_VEC", "BMODEL_CARTESIAN", "BMODEL"}
GRID_BUILD = False
INTRINSICS = {"MOD", "SQRT", "ABS", "COS", "SIN", "ACOS", "ATAN2", "MAX", "SUM", "DOT_PRODUCT", "MAXVAL", "MINVAL", "ALLOCATED", "MATMUL", "TRANSPOSE", "EXP"}
CTYPE = {"REAL*8": "double", "REAL(KIND=DP)": "double", "REAL(KIND=SP)": "float", "COMPLEX(KIND=DP)": "cd", "INTEGER": "int", "LOGICAL": "bool", "CHARACTER": "char"}

TOKEN = re.compile(r"(\.(?:LT|GT|LE|GE|EQ|NE|AND|OR|NOT|TRUE|FALSE)\.)|(\*\*|\(/|/\)|==|/=|>=|<=|[<>])|([A-Z][A-Z0-9_]*)|([(),+\-*/=:])")
NUMBER = re.compile(r"(\d+\.?\d*(?:[DE][+-]?\d+)?|\.\d+(?:[DE][+-]?\d+)?)(_DP|_SP|_8)?")
LOGOPS = {".LT.": "<", ".GT.": ">", ".LE.": "<=", ".GE.": ">=", ".EQ.": "==", ".NE.": "!=", ".AND.": "&&", ".OR.": "||", ".NOT.": "!", "==": "==",
          "/=": "!=", ">=": ">=", "<=": "<=", "<": "<", ">": ">"}


def tokenize(s):
    toks, i = [], 0
    while i < len(s):
        if s[i].isdigit() or (s[i] == "." and i + 1 < len(s) and s[i + 1].isdigit()):
            m = NUMBER.match(s, i)
            txt, kind = m.group(1), m.group(2)
            # a dot that starts a logical operator does not belong to the number
            if txt.endswith(".") and re.match(r"\.(?:LT|GT|LE|GE|EQ|NE|AND|OR|NOT)\.", s[m.start() + len(txt) - 1:]):
                txt, kind = txt[:-1], None
                toks.append(("NUM", txt, kind))
                i = m.start() + len(txt)
                continue
            toks.append(("NUM", txt, kind))
            i = m.end()
            continue
        if s[i] == "'":  # character literal: only in PRINT statements, which are dropped
            j = s.index("'", i + 1)
            toks.append(("STR", s[i:j + 1], None))
            i = j + 1
            continue
        m = TOKEN.match(s, i)
        if not m:
            raise ValueError("cannot tokenize %r at %r" % (s, s[i:]))
        toks.append(m.group(0))
        i = m.end()
    return toks


def number_from(text):
    sign = ""
    if text[0] in "+-":
        sign, text = text[0], text[1:]
    m = NUMBER.fullmatch(text)
    return sign + number_cpp(m.group(1), m.group(2))


def number_cpp(txt, kind):
    isreal = "." in txt or "E" in txt or "D" in txt
    if not isreal:
        return txt
    t = txt.replace("D", "e").replace("E", "e")
    if "." not in t.split("e")[0]:
        t = t.replace("e", ".e") if "e" in t else t + "."
    return t if (kind in ("_DP", "_8") or "D" in txt) else t + "f"  # a literal without a kind is single precision


class Sym:
    def __init__(self, ctype, rank=0, dims=None, alloc=False, param=None):
        self.ctype, self.rank, self.dims, self.alloc, self.param = ctype, rank, dims, alloc, param

    def cpp(self):
        if self.ctype == "char":
            return "const char*"
        return self.ctype if self.rank == 0 else {1: "Vec<%s>", 2: "Mat<%s>", 3: "Arr3<%s>", 4: "Arr4<%s>"}[self.rank] % self.ctype


class Unit:
    def __init__(self, kind, name, args):
        self.kind, self.name, self.args, self.syms, self.body = kind, name, args, {}, []


class Expr:
    """expression -> C++ text.  `scope` maps names to Sym (unit symbols over module symbols); `units` the procedures."""

    def __init__(self, toks, unit, scope, units):
        self.t, self.i, self.u, self.scope, self.units = toks, 0, unit, scope, units

    def peek(self, k=0):
        return self.t[self.i + k] if self.i + k < len(self.t) else None

    def take(self, want=None):
        tok = self.peek()
        if want is not None and tok != want:
            raise ValueError("expected %r, got %r in %r (unit %s)" % (want, tok, self.t, self.u.name))
        self.i += 1
        return tok

    def expr(self):
        a = self.and_()
        while self.peek() == ".OR.":
            self.take()
            a = "(%s || %s)" % (a, self.and_())
        return a

    def and_(self):
        a = self.not_()
        while self.peek() == ".AND.":
            self.take()
            a = "(%s && %s)" % (a, self.not_())
        return a

    def not_(self):
        if self.peek() == ".NOT.":
            self.take()
            return "(!%s)" % self.not_()
        return self.rel()

    def rel(self):
        a = self.add()
        if self.peek() in (".LT.", ".GT.", ".LE.", ".GE.", ".EQ.", ".NE.", "==", "/=", ">=", "<=", "<", ">"):
            op = LOGOPS[self.take()]
            a = "(%s %s %s)" % (a, op, self.add())
        return a

    def add(self):
        if self.peek() in ("+", "-"):
            op = self.take()
            a = "(%s%s)" % (op, self.mul())
        else:
            a = self.mul()
        while self.peek() in ("+", "-"):
            op = self.take()
            a = "(%s %s %s)" % (a, op, self.mul())
        return a

    def mul(self):
        a = self.pow_()
        while self.peek() in ("*", "/"):
            op = self.take()
            a = "(%s %s %s)" % (a, op, self.pow_())
        return a

    def pow_(self):
        base = self.primary()
        if self.peek() == "**":
            self.take()
            tok = self.peek()
            if isinstance(tok, tuple) and tok[0] == "NUM" and re.fullmatch(r"\d+", tok[1]):
                self.take()
                return "F_powc<%s>(%s)" % (tok[1], base)
            return "F_pow(%s, %s)" % (base, self.pow_())
        return base

    def subscripts(self):
        """after '(' : list of ('i', expr) | ('r', lo or None, hi or None); consumes ')'"""
        out = []
        while True:
            if self.peek() == ":":
                self.take()
                hi = None if self.peek() in (",", ")") else self.expr()
                out.append(("r", None, hi))
            else:
                e = self.expr()
                if self.peek() == ":":
                    self.take()
                    hi = None if self.peek() in (",", ")") else self.expr()
                    out.append(("r", e, hi))
                else:
                    out.append(("i", e))
            if self.peek() == ",":
                self.take()
                continue
            self.take(")")
            return out

    def array_ref(self, name, sym, lhs=False):
        subs = self.subscripts()
        v = "v_" + name
        if sym.rank == 1:
            s = subs[0]
            if s[0] == "i":
                return "%s(%s)" % (v, s[1])
            lo = s[1] or "1"
            hi = s[2] or "%s.size()" % v
            return "%s.%s(%s, %s)" % (v, "sec" if lhs else "rsec", lo, hi)
        if sym.rank == 3:
            if all(x[0] == "i" for x in subs):
                return "%s(%s)" % (v, ", ".join(x[1] for x in subs))
            ext = []  # a section; a scalar subscript keeps its dimension with extent 1 (every operand of one statement has it in the same place)
            for d, x in enumerate(subs):
                ext += [x[1], x[1]] if x[0] == "i" else [x[1] or "1", x[2] or "%s.n%d" % (v, d + 1)]
            return "%s.sec3(%s)" % (v, ", ".join(ext))
        if sym.rank == 4:
            if subs[0][0] == "i" and all(x[0] == "r" and x[1] is None and x[2] is None for x in subs[1:]):
                return "%s.slice1(%s)" % (v, subs[0][1])
            if all(x[0] == "i" for x in subs):
                return "%s(%s)" % (v, ", ".join(x[1] for x in subs))
            raise ValueError("rank-4 reference %s%r" % (name, subs))
        a, b = subs
        if a[0] == "i" and b[0] == "i":
            return "%s(%s, %s)" % (v, a[1], b[1])
        if a[0] == "r" and b[0] == "i" and a[1] is None and a[2] is None:
            return "%s.%s(%s)" % (v, "col" if lhs else "rcol", b[1])
        if a[0] == "i" and b[0] == "r" and not lhs:
            return "%s.row(%s)" % (v, a[1])
        if a[0] == "r" and b[0] == "r" and lhs:
            return "%s.block(%s, %s, %s, %s)" % (v, a[1] or "1", a[2] or v + ".n1", b[1] or "1", b[2] or v + ".n2")
        raise ValueError("array reference %s%r (lhs=%s)" % (name, subs, lhs))

    def call_args(self, callee):
        """actual arguments of a user procedure (after '('), dummy procedures dropped, temporaries for non-variables"""
        actual = []
        if self.peek() == ")":
            self.take()
        else:
            while True:
                start = self.i
                tok = self.peek()
                plain = isinstance(tok, str) and re.fullmatch(r"[A-Z][A-Z0-9_]*", tok) and self.peek(1) in (",", ")")
                if plain:
                    self.take()
                    actual.append(("var", tok))
                else:
                    self.i = start
                    actual.append(("expr", self.expr()))
                if self.peek() == ",":
                    self.take()
                    continue
                self.take(")")
                break
        dummies = callee.args
        if len(actual) != len(dummies):
            raise ValueError("%s calls %s with %d arguments, it has %d" % (self.u.name, callee.name, len(actual), len(dummies)))
        out = []
        for (kind, val), d in zip(actual, dummies):
            if d in DUMMY_PROCS:
                continue
            want = callee.syms[d].cpp()
            if kind == "var" and val in self.scope and self.scope[val].cpp() == want and self.scope[val].param is None:
                out.append("v_" + val)
            else:
                out.append("F_tmp<%s>(%s)" % (want, "v_" + val if kind == "var" else val))
        return ", ".join(out)

    def primary(self):
        tok = self.take()
        if tok == "(":
            e = self.expr()
            self.take(")")
            return "(%s)" % e
        if tok == "(/" and self.peek() == "(" and self.peek(2) == "," and self.peek(1) == self.peek(3) and self.peek(4) == "=":
            self.take(); self.take(); self.take(); self.take(); self.take()  # (/ (IND, IND = lo, hi) /): the integers lo..hi
            lo = self.expr(); self.take(","); hi = self.expr(); self.take(")"); self.take("/)")
            return "F90::iota(%s, %s)" % (lo, hi)
        if tok == "(/":
            items = []
            while True:
                items.append(self.expr())
                if self.peek() == ",":
                    self.take()
                    continue
                self.take("/)")
                break
            return "F_cat(%s)" % ", ".join(items)
        if isinstance(tok, tuple):
            return number_cpp(tok[1], tok[2])
        if tok in (".TRUE.", ".FALSE."):
            return "true" if tok == ".TRUE." else "false"
        name = tok
        if self.peek() != "(":
            if name in self.scope or name == self.u.name:
                return "v_" + name
            raise ValueError("unknown name %s in %s" % (name, self.u.name))
        self.take("(")
        if name in self.scope and self.scope[name].rank > 0:
            return self.array_ref(name, self.scope[name])
        if name == self.u.name and name in self.u.syms and self.u.syms[name].rank > 0:
            return self.array_ref(name, self.u.syms[name])
        if name == "REAL":
            e = self.expr()
            dp = False
            if self.peek() == ",":
                self.take(); self.take("KIND"); self.take("="); self.take(); dp = True
            self.take(")")
            return "F90::%s(%s)" % ("real" if dp else "real_default", e)
        if name == "CMPLX":
            a = self.expr(); self.take(","); b = self.expr()
            dp = False
            if self.peek() == ",":
                self.take(); self.take("KIND"); self.take("="); self.take(); dp = True
            self.take(")")
            return "F90::cmplx_%s(%s, %s)" % ("dp" if dp else "sp", a, b)
        if name in ("IMAG", "AIMAG"):
            e = self.expr(); self.take(")")
            return "F90::imag(%s)" % e
        if name == "SIZE":
            e = self.expr()
            if self.peek() == ",":
                self.take(); d = self.expr(); self.take(")")
                return "F90::size(%s, %s)" % (e, d)
            self.take(")")
            return "F90::size(%s)" % e
        if name == "RESHAPE":
            e = self.expr(); self.take(","); self.take("(/"); a = self.expr(); self.take(","); b = self.expr(); self.take("/)"); self.take(")")
            return "F90::reshape(%s, %s, %s)" % (e, a, b)
        if name in INTRINSICS:
            args = []
            while True:
                args.append(self.expr())
                if self.peek() == ",":
                    self.take()
                    continue
                self.take(")")
                break
            return "F90::%s(%s)" % (name.lower(), ", ".join(args))
        if name == "FUNCSTOPCONDITIONS":
            name = "RAYTRACER_STOPCONDITIONS"  # the driver passes raytracer_stopconditions (raytracer_driver.f95)
        if name not in self.units:
            raise ValueError("unknown function %s in %s" % (name, self.u.name))
        return "f_%s(%s)" % (name, self.call_args(self.units[name]))


def tr(text, unit, scope, units):
    p = Expr(tokenize(text), unit, scope, units)
    e = p.expr()
    if p.peek() is not None:
        raise ValueError("trailing tokens in %r: %r" % (text, p.t[p.i:]))
    return e


HEAD = re.compile(r"^(SUBROUTINE|FUNCTION)([A-Z][A-Z0-9_]*)\((.*)\)$")


def parse_decl(s, syms, prefix=""):
    spec, rest = s.split("::", 1)
    parts = split_top(spec)
    base = parts[0]
    if base.startswith("CHARACTER("):
        base = "CHARACTER"
    if base not in CTYPE:
        raise ValueError("declaration %r" % s)
    attrs = set(p.split("(")[0] for p in parts[1:])
    for item in split_top(rest):
        init = None
        if top_level_eq(item) >= 0:
            k = top_level_eq(item)
            item, init = item[:k], item[k + 1:]
        m = re.fullmatch(r"([A-Z][A-Z0-9_]*)(?:\((.+)\))?", item)
        dims = split_top(m.group(2)) if m.group(2) else []
        syms[prefix + m.group(1)] = Sym(CTYPE[base], len(dims), dims, "ALLOCATABLE" in attrs, init if "PARAMETER" in attrs else None)


DATA_ROWS = {}


def parse(stmts):
    """-> (module-level symbols, units)"""
    glob, units, cur, skip, prefix = {}, {}, None, False, ""
    for lab, s in stmts:
        s = s.rstrip(";").replace("DATAP%P%", "NGO_").replace("DAT%", "NGO_")
        if s.startswith("TYPE::"):  # components of the adapter's state type become module variables NGO_<component>
            prefix = "NGO_" if s in ("TYPE::NGOSTATEDATA", "TYPE::INTERPSTATEDATA") else "SKIP_"
            continue
        if s.startswith("ENDTYPE"):
            prefix = ""
            continue
        if s.startswith("COMMON/") or (s.startswith("TYPE(") and cur is None):
            continue
        if s.startswith("INTERFACE"):
            skip = True
            continue
        if s.startswith("ENDINTERFACE"):
            skip = False
            continue
        if skip:
            continue
        if re.fullmatch(r"(MODULE[A-Z0-9_]+|ENDMODULE.*|CONTAINS|IMPLICITNONE|USE[A-Z0-9_,:]+)", s) and top_level_eq(s) < 0:
            continue
        m = HEAD.match(s)
        if m and cur is None and top_level_eq(s) < 0:
            cur = Unit(m.group(1), m.group(2), [a for a in m.group(3).split(",") if a])
            continue
        if s.startswith("ENDFUNCTION") or s.startswith("ENDSUBROUTINE"):
            if cur.name in KEEP_UNITS or (GRID_BUILD and cur.name == "SETUP"):
                units[cur.name] = cur
            cur = None
            continue
        m2 = re.fullmatch(r"DATA([A-Z][A-Z0-9_]*)\((\d+),:\)/(.+)/", s)
        if m2 and cur is None:
            DATA_ROWS.setdefault(m2.group(1), {})[int(m2.group(2))] = split_top(m2.group(3))
            continue
        if "::" in s and (s.split("::")[0].split(",")[0] in CTYPE or s.startswith("CHARACTER(")):
            parse_decl(s, glob if cur is None else cur.syms, prefix if cur is None else "")
            continue
        if cur is not None:
            cur.body.append(s)
    return glob, units


def lhs_cpp(lhs, unit, scope, units):
    m = re.fullmatch(r"([A-Z][A-Z0-9_]*)(\(.*\))?", lhs)
    name = m.group(1)
    sym = scope.get(name)
    if sym is None:
        raise ValueError("assignment to unknown %s in %s" % (name, unit.name))
    if not m.group(2):
        return "v_" + name, sym, True
    p = Expr(tokenize(lhs), unit, scope, units)
    p.take(); p.take("(")
    return p.array_ref(name, sym, lhs=True), sym, False


def statement(s, unit, scope, units):
    if s.startswith("PRINT*") or s.startswith("WRITE("):
        return []
    if s.startswith("IF("):
        depth, j = 0, 2
        while True:
            if s[j] == "(":
                depth += 1
            elif s[j] == ")":
                depth -= 1
                if depth == 0:
                    break
            j += 1
        cond, rest = tr(s[3:j], unit, scope, units), s[j + 1:]
        if rest == "THEN":
            return ["if (%s) {" % cond]
        return ["if (%s) { %s }" % (cond, " ".join(statement(rest, unit, scope, units)))]
    if s.startswith("ELSEIF(") and s.endswith(")THEN"):
        return ["} else if (%s) {" % tr(s[7:-5], unit, scope, units)]
    if s == "ELSE":
        return ["} else {"]
    if s in ("ENDIF", "ENDDO"):
        return ["}"]
    if s == "DO":
        return ["for (;;) {"]
    m = re.fullmatch(r"DO([A-Z][A-Z0-9_]*)=(.+)", s)
    if m and len(split_top(m.group(2))) >= 2 and m.group(1) in scope:
        lim = split_top(m.group(2))
        var = "v_" + m.group(1)
        return ["for (%s = %s; %s <= %s; ++%s) {" % (var, tr(lim[0], unit, scope, units), var, tr(lim[1], unit, scope, units), var)]
    m = re.fullmatch(r"SELECTCASE\((.+)\)", s)
    if m:
        return ["switch (%s) {" % tr(m.group(1), unit, scope, units)]
    m = re.fullmatch(r"CASE\((\d+)\)", s)
    if m:
        return ["break; case %s:" % m.group(1)]
    if s == "CASEDEFAULT":
        return ["break; default:"]
    if s == "ENDSELECT":
        return ["}"]
    m = re.fullmatch(r"([A-Z][A-Z0-9_]*)=MAXLOC\(\(/\(IND,IND=1,([A-Z]+)\)/\),MASK=0\.0_8<=\(([A-Z]+)-([A-Z]+)\)\)", s)
    if m:  # the cell search of tricubic_interpolate_at (libtricubic.f95:834-839)
        return ["v_%s(1) = F90::maxloc_iota_mask(v_%s, v_%s, v_%s);" % m.groups()]
    m = re.fullmatch(r'OPEN\(UNIT=([A-Z0-9_]+),FILE=([A-Z0-9_]+),STATUS="OLD"\)', s.upper())
    if m:
        return ["F_open(v_%s, v_%s);" % m.groups()]
    m = re.fullmatch(r"CLOSE\(([A-Z0-9_]+)\)", s)
    if m:
        return ["F_close(v_%s);" % m.group(1)]
    m = re.fullmatch(r"READ\(([A-Z0-9_]+),\*\),?(.+)", s)
    if m:  # list-directed READ: one statement = one F_Read, which starts at a record and ends by skipping the rest of the last one
        def items(text):
            out = []
            for it in split_top(text):
                if it.startswith("(") and it.endswith(")") and top_level_eq(it[1:-1]) >= 0:
                    parts = split_top(it[1:-1])
                    k = next(i for i, q in enumerate(parts) if top_level_eq(q) >= 0)
                    var, lo = parts[k].split("=", 1)
                    v = "v_" + var
                    out.append("for (%s = %s; %s <= %s; ++%s) { %s }" % (v, tr(lo, unit, scope, units), v, tr(parts[k + 1], unit, scope, units), v,
                                                                         " ".join(items(",".join(parts[:k])))))
                else:
                    out.append("rd.item(%s);" % lhs_cpp(it, unit, scope, units)[0])
            return out
        return ["{ F_Read rd(v_%s); %s }" % (m.group(1), " ".join(items(m.group(2))))]
    if s == "EXIT":
        return ["break;"]
    if s == "CYCLE":
        return ["continue;"]
    if s == "RETURN":
        return ["return%s;" % (" v_" + unit.name if unit.kind == "FUNCTION" else "")]
    m = re.fullmatch(r"ALLOCATE\(([A-Z][A-Z0-9_]*)\((.+)\)\)", s)
    if m:
        sym = scope[m.group(1)]
        dims = [tr(d, unit, scope, units) for d in split_top(m.group(2))]
        return ["v_%s = %s(%s);" % (m.group(1), sym.cpp(), ", ".join(dims))]
    m = re.fullmatch(r"CALLMOVE_ALLOC\(([A-Z0-9_]+),([A-Z0-9_]+)\)", s)
    if m:
        return ["F_move_alloc(v_%s, v_%s);" % (m.group(1), m.group(2))]
    if s == "CALLDENS":
        return ["RT_dens();"]
    if s.startswith("TYPE(") or s.startswith("DATAP=") or s.startswith("COMMON/"):
        return []
    m = re.fullmatch(r"CALL([A-Z][A-Z0-9_]*)\((.*)\)", s)
    if m and m.group(1) in HOOK_SUBS:
        return ["%s(%s);" % (HOOK_SUBS[m.group(1)], ", ".join(tr(a, unit, scope, units) for a in split_top(m.group(2))))]
    if m:
        name = m.group(1)
        p = Expr(tokenize("(" + m.group(2) + ")"), unit, scope, units)
        p.take("(")
        if name == "FUNCPLASMAPARAMS":
            args = []
            for k in range(7):
                tok = p.peek()
                plain = isinstance(tok, str) and re.fullmatch(r"[A-Z][A-Z0-9_]*", tok) and p.peek(1) in (",", ")")
                if plain:
                    p.take(); args.append("v_" + tok)
                else:
                    args.append("F_tmp<Vec<double>>(%s)" % p.expr())
                p.take("," if k < 6 else ")")
            return ["RT_plasma(%s);" % ", ".join(args[:6])]
        if name == "CSVD":
            names = [t for t in p.t[1:-1] if t != ","]
            return ["RT_csvd(%s);" % ", ".join("v_" + n for n in names)]
        if name not in units:
            raise ValueError("unknown subroutine %s in %s" % (name, unit.name))
        return ["f_%s(%s);" % (name, p.call_args(units[name]))]
    k = top_level_eq(s)
    if k >= 0:
        target, sym, whole = lhs_cpp(s[:k], unit, scope, units)
        rhs = tr(s[k + 1:], unit, scope, units)
        if sym.rank == 0 and sym.ctype == "double":
            return ["%s = F_toreal(%s);" % (target, rhs)]
        if whole and sym.rank == 1 and sym.ctype == "cd":
            return ["%s = F_tocplx(%s);" % (target, rhs)]
        if sym.rank == 3 and (whole or ".sec3(" in target):
            return ["F_assign3(%s, %s);" % (target, rhs)]
        return ["%s = %s;" % (target, rhs)]
    raise ValueError("cannot translate %r in %s" % (s, unit.name))


def emit(glob, units):
    out, protos = [], []
    consts = []
    for name, sym in glob.items():  # named constants of the modules, in source order
        if name in DATA_ROWS:  # a module matrix initialised row by row in DATA statements
            rows = DATA_ROWS[name]
            fill = " ".join("{ const double r[] = {%s}; for (int j = 0; j < %d; ++j) m(%d, j + 1) = r[j]; }" % (", ".join(number_from(v) for v in vals), len(vals), r)
                            for r, vals in sorted(rows.items()))
            consts.append("static Mat<double> F_data_%s() { Mat<double> m(%s); %s return m; }\nstatic const Mat<double> v_%s = F_data_%s();" % (
                name, ", ".join(sym.dims), fill, name, name))
            continue
        if sym.param is None:
            if name not in ("SP", "DP"):
                dims = ", ".join(sym.dims or [])
                fixed = sym.rank in (1, 2) and ":" not in (sym.dims or [])
                consts.append("static thread_local %s v_%s%s;" % (sym.cpp(), name, "(%s)" % dims if fixed else ("" if sym.rank else " = 0")))
            continue
        dummy = Unit("MODULE", "", [])
        consts.append("static const %s v_%s = %s;" % (sym.cpp(), name, tr(sym.param, dummy, glob, units)))
    for u in units.values():
        scope = dict(glob)
        scope.update(u.syms)
        params = ", ".join("%s& v_%s" % (u.syms[a].cpp(), a) for a in u.args if a not in DUMMY_PROCS)
        rt = "void" if u.kind == "SUBROUTINE" else u.syms[u.name].cpp()
        decl = []
        for n, sym in u.syms.items():
            if n in u.args or n in DUMMY_PROCS:
                continue
            dims = [d for d in (sym.dims or [])]
            if sym.param is not None:  # a local named constant
                decl.append("const %s v_%s = %s;" % (sym.cpp(), n, tr(sym.param, u, scope, units)))
            elif sym.rank and not sym.alloc and ":" not in dims:
                decl.append("%s v_%s(%s);" % (sym.cpp(), n, ", ".join(tr(d, u, scope, units) for d in dims)))
            elif sym.rank:
                decl.append("%s v_%s;" % (sym.cpp(), n))
            else:
                decl.append("%s v_%s = %s;" % (sym.cpp(), n, "false" if sym.ctype == "bool" else "0"))  # -finit-local-zero (fortran/Makefile)
        body = []
        for s in u.body:
            body += statement(s, u, scope, units)
        tail = ["return v_%s;" % u.name] if u.kind == "FUNCTION" else []
        protos.append("%s f_%s(%s);" % (rt, u.name, params))
        out.append("%s f_%s(%s) {\n  %s\n  %s\n}\n" % (rt, u.name, params, "\n  ".join(decl), "\n  ".join(body + tail)))
    return consts, protos, out


WRAPPERS = r"""
// ---- the Ngo adapter (fortran/ngo_dens_model_adapter.f95:63-207) calls into the FORTRAN 77 libraries of the reference; those are
// translated into their own shared objects (libngo_ref, libxform_ref, libgeopack_ref, libts05_ref): the harness hands in their entry points
typedef void (*xform_fn)(int, int, const double*, double*);
typedef void (*recalc_fn)(int, int, int, int, int, float*, float*);
typedef void (*igrf_fn)(float, float, float, float*);
typedef void (*t04_fn)(const float*, float, float, float, float, float*);
typedef void (*dens_fn)(double, double, double, double*);
static thread_local xform_fn g_sm2gsm = nullptr, g_gsm2sm = nullptr;
static thread_local recalc_fn g_recalc = nullptr;
static thread_local igrf_fn g_igrf = nullptr;
static thread_local t04_fn g_t04 = nullptr;
static thread_local dens_fn g_dens = nullptr;
#if defined(RT_HAVE_NGO_ADAPTER) || defined(RT_HAVE_GRID_ADAPTER)
static void RT_sm_to_gsm(const Vec<int>& itime, const Vec<double>& x, Vec<double>& out) { g_sm2gsm(itime(1), itime(2), x.d.data(), out.d.data()); }
static void RT_gsm_to_sm(const Vec<int>& itime, const Vec<double>& x, Vec<double>& out) { g_gsm2sm(itime(1), itime(2), x.d.data(), out.d.data()); }
static void RT_tsy_recalc(int year, int day, int hour, int min, int sec) {
  float g1[34], g2[315];
  g_recalc(year, day, hour, min, sec, g1, g2);
  v_PSI = g1[0];  // `real(kind=SP) :: PSI; COMMON /GEOPACK1/ PSI`: word 1 of the block
}
static void RT_igrf_gsm(float x, float y, float z, float& bx, float& by, float& bz) { float b[3]; g_igrf(x, y, z, b); bx = b[0]; by = b[1]; bz = b[2]; }
static void RT_t04_s(int iopt, const Vec<float>& parmod, float ps, float x, float y, float z, float& bx, float& by, float& bz) {
  float b[3]; (void)iopt; g_t04(parmod.d.data(), ps, x, y, z, b); bx = b[0]; by = b[1]; bz = b[2];
}
#endif
#if defined(RT_HAVE_NGO_ADAPTER)
static void RT_dens() { double a[4]; g_dens(v_Z(1), v_Z(2), v_LATITU, a); for (int i = 1; i <= 4; ++i) v_ANI(i) = a[i - 1]; }
#endif

// ---- hooks of the test harness and C entry points (ctypes)
typedef void (*plasma_cb_t)(const double* x, double* qs, double* Ns, double* ms, double* nus, double* B0, int* nspec);
typedef void (*csvd_cb_t)(const double* m_re_im /*[3][3][2] column-major*/, double* vt_re_im /*[3][3][2]*/);
static thread_local plasma_cb_t g_plasma = nullptr;
static thread_local csvd_cb_t g_csvd = nullptr;
static thread_local long g_plasma_calls = 0;
static void RT_plasma(Vec<double>& x, Vec<double>& qs, Vec<double>& Ns, Vec<double>& ms, Vec<double>& nus, Vec<double>& B0) {
  double q[8], n[8], m[8], nu[8], b[3]; int ns = 0;
  g_plasma(x.d.data(), q, n, m, nu, b, &ns);
  ++g_plasma_calls;
  if (!qs.allocated) qs = Vec<double>(ns);   // the adapters allocate on first use (ngo_dens_model_adapter.f95:91-102)
  if (!Ns.allocated) Ns = Vec<double>(ns);
  if (!ms.allocated) ms = Vec<double>(ns);
  if (!nus.allocated) nus = Vec<double>(ns);
  for (int i = 1; i <= ns; ++i) { qs(i) = q[i - 1]; Ns(i) = n[i - 1]; ms(i) = m[i - 1]; nus(i) = nu[i - 1]; }
  if (B0.size() != 3) B0 = Vec<double>(3);
  for (int i = 1; i <= 3; ++i) B0(i) = b[i - 1];
}
// csvd (blas.f95:180-221 -> LAPACK zgesvd, not translated): only VT is used by the caller
static void RT_csvd(Mat<cd>& M, Mat<cd>& U, Vec<double>& svd, Mat<cd>& VT) {
  double in[18], out[18];
  for (int j = 1; j <= 3; ++j) for (int i = 1; i <= 3; ++i) { in[2 * ((i - 1) + 3 * (j - 1))] = M(i, j).real(); in[2 * ((i - 1) + 3 * (j - 1)) + 1] = M(i, j).imag(); }
  g_csvd(in, out);
  for (int j = 1; j <= 3; ++j) for (int i = 1; i <= 3; ++i) VT(i, j) = cd(out[2 * ((i - 1) + 3 * (j - 1))], out[2 * ((i - 1) + 3 * (j - 1)) + 1]);
  (void)U; (void)svd;
}
"""

ENTRY = r"""
extern "C" {
void rt_ref_set_hooks(plasma_cb_t p, csvd_cb_t c) { g_plasma = p; g_csvd = c; }
long rt_ref_plasma_calls() { return g_plasma_calls; }
double rt_ref_constant(int which) { const double c[6] = {v_EPS0, v_PI, v_MU0, v_C, v_R_E, v_R2D}; return c[which]; }
double rt_ref_dispersion_relation(const double* n, double w, int nspec, const double* qs, const double* Ns, const double* ms, const double* nus, const double* B0) {
  Vec<double> vn(3), vq(nspec), vN(nspec), vm(nspec), vnu(nspec), vb(3);
  for (int i = 1; i <= 3; ++i) { vn(i) = n[i - 1]; vb(i) = B0[i - 1]; }
  for (int i = 1; i <= nspec; ++i) { vq(i) = qs[i - 1]; vN(i) = Ns[i - 1]; vm(i) = ms[i - 1]; vnu(i) = nus[i - 1]; }
  return f_DISPERSION_RELATION(vn, w, vq, vN, vm, vnu, vb);
}
void rt_ref_stix(double w, int nspec, const double* qs, const double* Ns, const double* ms, const double* nus, double B0mag, double* sdprl) {
  Vec<double> vq(nspec), vN(nspec), vm(nspec), vnu(nspec);
  for (int i = 1; i <= nspec; ++i) { vq(i) = qs[i - 1]; vN(i) = Ns[i - 1]; vm(i) = ms[i - 1]; vnu(i) = nus[i - 1]; }
  f_STIX_PARAMETERS(w, vq, vN, vm, vnu, B0mag, sdprl[0], sdprl[1], sdprl[2], sdprl[3], sdprl[4]);
}
void rt_ref_evalrhs(const double* args7, double del, double* out7) {
  Vec<double> a(7); for (int i = 1; i <= 7; ++i) a(i) = args7[i - 1];
  double t = 0.0;
  Vec<double> r = f_RAYTRACER_EVALRHS(t, a, del);
  for (int i = 1; i <= 7; ++i) out7[i - 1] = r(i);
}
void rt_ref_solve(const double* k, double w, const double* x, double* k1k2 /* re, im, re, im */) {
  Vec<double> vk(3), vx(3); for (int i = 1; i <= 3; ++i) { vk(i) = k[i - 1]; vx(i) = x[i - 1]; }
  cd k1, k2;
  f_SOLVE_DISPERSION_RELATION(vk, w, vx, k1, k2);
  k1k2[0] = k1.real(); k1k2[1] = k1.imag(); k1k2[2] = k2.real(); k1k2[3] = k2.imag();
}
int rt_ref_is_right_handed(double n2, double phi, double S, double D, double P) { return f_IS_RIGHT_HANDED(n2, phi, S, D, P) ? 1 : 0; }
void rt_ref_bmodel_cartesian(const double* x, double* b) {
  Vec<double> vx(3); for (int i = 1; i <= 3; ++i) vx(i) = x[i - 1];
  Vec<double> r = f_BMODEL_CARTESIAN(vx);
  for (int i = 1; i <= 3; ++i) b[i - 1] = r(i);
}
void rt_ref_cartesian_to_spherical(const double* x, double* p) {
  Vec<double> vx(3); for (int i = 1; i <= 3; ++i) vx(i) = x[i - 1];
  Vec<double> r = f_CARTESIAN_TO_SPHERICAL(vx);
  for (int i = 1; i <= 3; ++i) p[i - 1] = r(i);
}
// raytracer_run: rows out as [n][20] = t, pos3, vprel3, vgrel3, n3, B03, Ns4 (the layout of the oracle's and the product's rows)
long rt_ref_run(const double* pos0, double* dir0, double w0, double dt0, double dtmax, double maxerr, int maxsteps, double minalt, int root, double tmax,
                int fixedstep, double del, double* rows, long max_rows, int* stopcond_out) {
  Mat<double> pos, vprel, vgrel, n, B0, qs, ms, Ns, nus; Vec<double> time;
  Vec<double> p0(3), d0(3); for (int i = 1; i <= 3; ++i) { p0(i) = pos0[i - 1]; d0(i) = dir0[i - 1]; }
  int stopcond = 0;
  f_RAYTRACER_RUN(pos, time, vprel, vgrel, n, B0, qs, ms, Ns, nus, stopcond, p0, d0, w0, dt0, dtmax, maxerr, maxsteps, minalt, root, tmax, fixedstep, del);
  for (int i = 1; i <= 3; ++i) dir0[i - 1] = d0(i);
  *stopcond_out = stopcond;
  const long nr = time.size();
  for (long r = 1; r <= nr && r <= max_rows; ++r) {
    double* o = rows + 20 * (r - 1);
    o[0] = time((int)r);
    for (int i = 1; i <= 3; ++i) { o[i] = pos(i, (int)r); o[3 + i] = vprel(i, (int)r); o[6 + i] = vgrel(i, (int)r); o[9 + i] = n(i, (int)r); o[12 + i] = B0(i, (int)r); }
    for (int i = 1; i <= 4; ++i) o[15 + i] = i <= Ns.n1 ? Ns(i, (int)r) : 0.0;
  }
  return nr;
}
#if defined(RT_HAVE_NGO_ADAPTER)
// funcPlasmaParams of the Ngo adapter, translated: state = the adapter's ngoStateData, r0 = module ngo_dens_model's R0 after readinput
void rt_ref_ngo_adapter(void* sm2gsm, void* gsm2sm, void* recalc, void* igrf, void* t04, void* dens, double r0, int yearday, int msec, int use_igrf,
                        int use_tsyganenko, const double* parmod) {
  g_sm2gsm = (xform_fn)sm2gsm; g_gsm2sm = (xform_fn)gsm2sm; g_recalc = (recalc_fn)recalc; g_igrf = (igrf_fn)igrf; g_t04 = (t04_fn)t04; g_dens = (dens_fn)dens;
  v_R0 = r0; v_NGO_ITIME(1) = yearday; v_NGO_ITIME(2) = msec; v_NGO_USE_IGRF = use_igrf; v_NGO_USE_TSYGANENKO = use_tsyganenko;
  v_NGO_PDYN = parmod[0]; v_NGO_DST = parmod[1]; v_NGO_BYIMF = parmod[2]; v_NGO_BZIMF = parmod[3];
  v_NGO_W1 = parmod[4]; v_NGO_W2 = parmod[5]; v_NGO_W3 = parmod[6]; v_NGO_W4 = parmod[7]; v_NGO_W5 = parmod[8]; v_NGO_W6 = parmod[9];
}
#endif
#if defined(RT_HAVE_GRID_ADAPTER)
// funcPlasmaParams of the interp adapter (interp_dens_model_adapter.f95:148-269), translated with libtricubic.f95: the state
// arrays (interpStateData: f and its seven derivative arrays (nspec, nx, ny, nz), node coordinates, qs, ms) are the harness's
void rt_ref_grid_adapter(void* sm2gsm, void* gsm2sm, void* recalc, void* igrf, void* t04, int yearday, int msec, int use_igrf, int use_tsyganenko,
                         const double* parmod, const int* dims /* nspec nx ny nz */, const double* del3, const double* qs, const double* ms,
                         const double* x, const double* y, const double* z, double* const* arrays /* f dfdx dfdy dfdz d2fdxdy d2fdxdz d2fdydz d3fdxdydz */) {
  g_sm2gsm = (xform_fn)sm2gsm; g_gsm2sm = (xform_fn)gsm2sm; g_recalc = (recalc_fn)recalc; g_igrf = (igrf_fn)igrf; g_t04 = (t04_fn)t04;
  v_NGO_ITIME(1) = yearday; v_NGO_ITIME(2) = msec; v_NGO_USE_IGRF = use_igrf; v_NGO_USE_TSYGANENKO = use_tsyganenko;
  v_NGO_PDYN = parmod[0]; v_NGO_DST = parmod[1]; v_NGO_BYIMF = parmod[2]; v_NGO_BZIMF = parmod[3];
  v_NGO_W1 = parmod[4]; v_NGO_W2 = parmod[5]; v_NGO_W3 = parmod[6]; v_NGO_W4 = parmod[7]; v_NGO_W5 = parmod[8]; v_NGO_W6 = parmod[9];
  v_NGO_NSPEC = dims[0]; v_NGO_NX = dims[1]; v_NGO_NY = dims[2]; v_NGO_NZ = dims[3];
  v_NGO_DELX = del3[0]; v_NGO_DELY = del3[1]; v_NGO_DELZ = del3[2];
  v_NGO_QS = Vec<double>(dims[0]); v_NGO_MS = Vec<double>(dims[0]);
  for (int i = 1; i <= dims[0]; ++i) { v_NGO_QS(i) = qs[i - 1]; v_NGO_MS(i) = ms[i - 1]; }
  v_NGO_X = Vec<double>(dims[1]); v_NGO_Y = Vec<double>(dims[2]); v_NGO_Z = Vec<double>(dims[3]);
  for (int i = 1; i <= dims[1]; ++i) v_NGO_X(i) = x[i - 1];
  for (int i = 1; i <= dims[2]; ++i) v_NGO_Y(i) = y[i - 1];
  for (int i = 1; i <= dims[3]; ++i) v_NGO_Z(i) = z[i - 1];
  Arr4<double>* a[8] = {&v_NGO_F, &v_NGO_DFDX, &v_NGO_DFDY, &v_NGO_DFDZ, &v_NGO_D2FDXDY, &v_NGO_D2FDXDZ, &v_NGO_D2FDYDZ, &v_NGO_D3FDXDYDZ};
  for (int k = 0; k < 8; ++k) { a[k]->p = arrays[k]; a[k]->n1 = dims[0]; a[k]->n2 = dims[1]; a[k]->n3 = dims[2]; a[k]->n4 = dims[3]; }
}
// setup of the interp adapter (interp_dens_model_adapter.f95:52-134), translated: reads the grid file (list-directed READs),
// builds the node coordinates and either reads the seven derivative arrays or computes them by finite differences; the other
// members of the state (time, field switches, TS05 parameters) are the caller's, as in the driver (raytracer_driver.f95)
void rt_ref_grid_setup(const char* path, void* sm2gsm, void* gsm2sm, void* recalc, void* igrf, void* t04, int yearday, int msec, int use_igrf,
                       int use_tsyganenko, const double* parmod) {
  g_sm2gsm = (xform_fn)sm2gsm; g_gsm2sm = (xform_fn)gsm2sm; g_recalc = (recalc_fn)recalc; g_igrf = (igrf_fn)igrf; g_t04 = (t04_fn)t04;
  v_NGO_ITIME(1) = yearday; v_NGO_ITIME(2) = msec; v_NGO_USE_IGRF = use_igrf; v_NGO_USE_TSYGANENKO = use_tsyganenko;
  v_NGO_PDYN = parmod[0]; v_NGO_DST = parmod[1]; v_NGO_BYIMF = parmod[2]; v_NGO_BZIMF = parmod[3];
  v_NGO_W1 = parmod[4]; v_NGO_W2 = parmod[5]; v_NGO_W3 = parmod[6]; v_NGO_W4 = parmod[7]; v_NGO_W5 = parmod[8]; v_NGO_W6 = parmod[9];
  f_SETUP(path);
}
// what setup left in the state: dims = nspec nx ny nz; ptrs = qs ms x y z f dfdx dfdy dfdz d2fdxdy d2fdxdz d2fdydz d3fdxdydz
void rt_ref_grid_state(int* dims, double* del3, double** ptrs) {
  dims[0] = v_NGO_NSPEC; dims[1] = v_NGO_NX; dims[2] = v_NGO_NY; dims[3] = v_NGO_NZ;
  del3[0] = v_NGO_DELX; del3[1] = v_NGO_DELY; del3[2] = v_NGO_DELZ;
  ptrs[0] = &v_NGO_QS(1); ptrs[1] = &v_NGO_MS(1); ptrs[2] = &v_NGO_X(1); ptrs[3] = &v_NGO_Y(1); ptrs[4] = &v_NGO_Z(1);
  Arr4<double>* a[8] = {&v_NGO_F, &v_NGO_DFDX, &v_NGO_DFDY, &v_NGO_DFDZ, &v_NGO_D2FDXDY, &v_NGO_D2FDXDZ, &v_NGO_D2FDYDZ, &v_NGO_D3FDXDYDZ};
  for (int k = 0; k < 8; ++k) ptrs[5 + k] = a[k]->p;
}
// tricubic_compute_finite_difference_derivatives (libtricubic.f95:722-793), translated, called species by species on the
// sections (ind,:,:,:) as setup does (interp_dens_model_adapter.f95:120-131); arrays[0] = f is read, arrays[1..7] are written
void rt_ref_grid_fd(const int* dims, const double* del3, double* const* arrays) {
  Arr4<double> a[8];
  for (int k = 0; k < 8; ++k) { a[k].p = arrays[k]; a[k].n1 = dims[0]; a[k].n2 = dims[1]; a[k].n3 = dims[2]; a[k].n4 = dims[3]; }
  double dx = del3[0], dy = del3[1], dz = del3[2];
  for (int ind = 1; ind <= dims[0]; ++ind)
    f_TRICUBIC_COMPUTE_FINITE_DIFFERENCE_DERIVATIVES(F_tmp<Arr3<double>>(a[0].slice1(ind)), dx, dy, dz,
        F_tmp<Arr3<double>>(a[1].slice1(ind)), F_tmp<Arr3<double>>(a[2].slice1(ind)), F_tmp<Arr3<double>>(a[3].slice1(ind)), F_tmp<Arr3<double>>(a[4].slice1(ind)),
        F_tmp<Arr3<double>>(a[5].slice1(ind)), F_tmp<Arr3<double>>(a[6].slice1(ind)), F_tmp<Arr3<double>>(a[7].slice1(ind)));
}
#endif
#if defined(RT_HAVE_NGO_ADAPTER) || defined(RT_HAVE_GRID_ADAPTER)
static void ngo_adapter_cb(const double* x, double* qs, double* Ns, double* ms, double* nus, double* B0, int* nspec) {
  Vec<double> vx(3), q, n, m, nu, b(3);
  for (int i = 1; i <= 3; ++i) vx(i) = x[i - 1];
  f_FUNCPLASMAPARAMS(vx, q, n, m, nu, b);
  *nspec = q.size();
  for (int i = 1; i <= q.size(); ++i) { qs[i - 1] = q(i); Ns[i - 1] = n(i); ms[i - 1] = m(i); nus[i - 1] = nu(i); }
  for (int i = 1; i <= 3; ++i) B0[i - 1] = b(i);
}
// funcPlasmaParams of raytracer_run and friends = the translated adapter (csvd stays the hook given)
void rt_ref_use_ngo_adapter(csvd_cb_t c) { g_plasma = ngo_adapter_cb; g_csvd = c; }
void rt_ref_ngo_plasma(long n, const double* x, double* out /*[n][19]*/) {
  for (long i = 0; i < n; ++i) {
    double qs[8], Ns[8], ms[8], nus[8], b[3]; int ns = 0;
    ngo_adapter_cb(x + 3 * i, qs, Ns, ms, nus, b, &ns);
    double* o = out + 19 * i;
    for (int s = 0; s < 4; ++s) { o[s] = qs[s]; o[4 + s] = Ns[s]; o[8 + s] = ms[s]; o[12 + s] = nus[s]; }
    for (int c = 0; c < 3; ++c) o[16 + c] = b[c];
  }
}
#endif
}
"""


def main(dst, srcs):
    global GRID_BUILD
    GRID_BUILD = any("libtricubic" in one for one in srcs)
    stmts = []
    for one in srcs:
        stmts += read_statements_free(one)
    glob, units = parse(stmts)
    glob = {k: v for k, v in glob.items() if not k.startswith("SKIP_")}
    grid = "TRICUBIC_INTERPOLATE_AT" in units
    if "FUNCPLASMAPARAMS" in units and not grid:  # `use ngo_dens_model, only : readinput, dens, ani, z, L, R0, latitu`
        for name, sym in (("Z", Sym("double", 1, ["5"])), ("ANI", Sym("double", 1, ["4"])), ("L", Sym("double")), ("R0", Sym("double")), ("LATITU", Sym("double"))):
            glob.setdefault(name, sym)
    consts, protos, texts = emit(glob, units)
    with open(dst, "w") as f:
        f.write("// GENERATED by oracle/f90_to_cpp.py from the reference's Fortran 90 at build time -- never committed.\n#include \"f90rt.h\"\n")
        if "FUNCPLASMAPARAMS" in units:
            f.write("#define RT_HAVE_%s_ADAPTER 1\n" % ("GRID" if grid else "NGO"))
        f.write("\n".join(consts) + "\n" + WRAPPERS + "\n" + "\n".join(protos) + "\n\n" + "\n".join(texts) + ENTRY)


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2:])
