#!/usr/bin/env python
"""Mechanical translator of the fixed-form FORTRAN 77 subset used by the reference's tsyganenko/TS05_aka_TS04.for to C++.

TEST INFRASTRUCTURE (oracle/): there is no Fortran compiler in this environment, so the reference's own source of the
TS05 external field is translated statement by statement where it lies (/root/reference, read at build time only) and
compiled into oracle/_ref/ -- the translated text never enters the repository.  The translation keeps every operation,
its order and its type:
  * implicit typing (I-N integer, the rest REAL or what IMPLICIT says), REAL literals stay single precision (`4.7` is
    4.7f, also in DATA lists of REAL*8 arrays), mixed-mode arithmetic is left to C++'s identical promotion rules;
  * `x**n` with a literal integer n is the multiplication tree gfortran emits (gcc powi table), `x**i` with an integer
    variable is libgcc's __powidf2 loop, `x**y` is pow();
  * every local is static (the reference builds this directory with -fno-automatic), here thread_local as well;
  * arguments are passed by reference, arrays by pointer with 1-based indexing, COMMON blocks are global arrays
    matched by position, labels / GOTO map to C labels / goto.
Usage: f77_to_cpp.py in.for out.cpp
"""
import re
import sys

LOGOPS = {".LT.": "<", ".GT.": ">", ".LE.": "<=", ".GE.": ">=", ".EQ.": "==", ".NE.": "!=", ".AND.": "&&", ".OR.": "||", ".NOT.": "!"}
INTRINSICS = {"DSIN": "M_sin", "DCOS": "M_cos", "DSQRT": "M_sqrt", "SQRT": "M_sqrt", "DEXP": "M_exp", "DLOG": "M_log", "DATAN2": "M_atan2",
              "DABS": "M_abs", "DTAN": "M_tan", "TAN": "M_tan", "SIN": "M_sin", "COS": "M_cos", "EXP": "M_exp", "ABS": "M_abs", "ATAN2": "M_atan2",
              "ATAN": "M_atan", "ASIN": "M_asin", "FLOAT": "M_float", "DFLOAT": "M_dfloat", "DMOD": "M_mod", "AMOD": "M_mod", "MOD": "M_mod", "SIGN": "M_sign", "DSIGN": "M_sign"}


def read_statements(path):
    """-> list of (label, text) with comments dropped, continuation lines joined, blanks removed, upper case"""
    out = []
    for raw in open(path, errors="replace"):
        line = raw.rstrip("\n").rstrip("\r")
        if not line.strip() or line[0] in "cC*!":
            continue
        if "!" in line:
            line = line[:line.index("!")]
        line = line.expandtabs(8)[:72]  # columns 73+ are not part of the statement
        if not line.strip():
            continue
        if len(line) > 5 and line[:5].strip() == "" and line[5] not in " 0":
            out[-1][1] += line[6:72]
            continue
        out.append([line[:5].strip(), line[6:72]])
    return [(lab, re.sub(r"\s+", "", txt).upper()) for lab, txt in out]


def read_statements_free(path):
    """free source form (Fortran 90): `!` comments, `&` continuation, label = leading digits; same output as read_statements"""
    out, pending = [], None
    for raw in open(path, errors="replace"):
        line, q = "", None
        for ch in raw.rstrip("\n").rstrip("\r"):
            if q:
                q = None if ch == q else q
            elif ch in "'\"":
                q = ch
            elif ch == "!":
                break
            line += ch
        line = line.strip()
        if not line:
            continue
        if pending is not None:
            line = pending + line.lstrip("&")
            pending = None
        if line.endswith("&"):
            pending = line[:-1]
            continue
        m = re.match(r"^(\d+)\s+(.*)$", line)
        lab, txt = (m.group(1), m.group(2)) if m else ("", line)
        out.append((lab, re.sub(r"\s+", "", txt).upper()))
    return out


TOKEN = re.compile(r"(\.(?:LT|GT|LE|GE|EQ|NE|AND|OR|NOT|TRUE|FALSE)\.)|(\*\*)|([A-Z][A-Z0-9_]*)|([(),+\-*/=])")


def tokenize(s):
    toks, i = [], 0
    while i < len(s):
        m = TOKEN.match(s, i)
        if m and not (s[i] == "." and not m.group(1)):
            toks.append(m.group(0))
            i = m.end()
            continue
        # number: digits [. digits] [exponent], the dot is not taken if it starts a logical operator
        j = i
        while j < len(s) and s[j].isdigit():
            j += 1
        isreal = False
        if j < len(s) and s[j] == "." and not re.match(r"\.(?:LT|GT|LE|GE|EQ|NE|AND|OR|NOT)\.", s[j:]):
            isreal = True
            j += 1
            while j < len(s) and s[j].isdigit():
                j += 1
        m2 = re.match(r"[DE][+-]?\d+", s[j:])
        if m2 and (isreal or j > i):
            j += m2.end()
        if j == i:
            raise ValueError("cannot tokenize %r at %r" % (s, s[i:]))
        text = s[i:j]
        if s[j:j + 2] == "_8" and not s[j + 2:j + 3].isalnum():  # kind suffix: a REAL*8 literal
            text = text.replace("E", "D") if "E" in text else text + "D0"
            j += 2
        toks.append(("NUM", text))
        i = j
    return toks


def number_to_cpp(txt):
    if "D" in txt:
        return txt.replace("D", "e") if "." in txt.split("D")[0] else txt.replace("D", ".e")
    if "." in txt or "E" in txt:
        t = txt.replace("E", "e")
        if "." not in t.split("e")[0]:
            t = t.replace("e", ".e")
        return t + "f"  # REAL*4 literal
    return txt  # integer


class Unit:
    """one SUBROUTINE / FUNCTION"""

    def __init__(self, kind, name, args, rtype):
        self.kind, self.name, self.args, self.rtype = kind, name, args, rtype
        self.implicit_double = False
        self.implicit = {}       # letter -> type (IMPLICIT REAL(KIND=DP) (A-H))
        self.module_vars = set() # names that live in the enclosing module (free form)
        self.types = {}    # explicit types
        self.dims = {}     # array name -> dimension text (1-D only)
        self.commons = []  # (block, [names])
        self.data = []     # (name list, value list)
        self.body = []     # (label, text) executable statements
        self.names = set()

    def type_of(self, name):
        if name in self.types:
            return self.types[name]
        if name[0] in self.implicit:
            return self.implicit[name[0]]
        if name[0] in "IJKLMN":
            return "int"
        return "double" if self.implicit_double else "float"


class ExprParser:
    """Fortran expression -> C++ text; precedence: .OR. < .AND. < .NOT. < relational < + - < * / < **"""

    def __init__(self, toks, unit, units):
        self.t, self.i, self.u, self.units = toks, 0, unit, units

    def peek(self):
        return self.t[self.i] if self.i < len(self.t) else None

    def take(self, want=None):
        tok = self.peek()
        if want is not None and tok != want:
            raise ValueError("expected %r, got %r in %r" % (want, tok, self.t))
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
        if self.peek() in (".LT.", ".GT.", ".LE.", ".GE.", ".EQ.", ".NE."):
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
            # right associative; the exponent may carry a sign only inside parentheses
            tok = self.peek()
            if isinstance(tok, tuple) and re.fullmatch(r"\d+", tok[1]) and self.t[self.i + 1:self.i + 2] != ["**"]:
                self.take()
                return "F_powc<%s>(%s)" % (tok[1], base)
            return "F_pow(%s, %s)" % (base, self.pow_())
        return base

    def primary(self):
        tok = self.take()
        if tok == "(":
            e = self.expr()
            self.take(")")
            return "(%s)" % e
        if isinstance(tok, tuple):
            return number_to_cpp(tok[1])
        if tok in (".TRUE.", ".FALSE."):
            return "true" if tok == ".TRUE." else "false"
        if not re.fullmatch(r"[A-Z][A-Z0-9_]*", tok):
            raise ValueError("unexpected token %r in %r" % (tok, self.t))
        name = tok
        if self.peek() != "(":
            self.u.names.add(name)
            return "v_" + name
        self.take("(")
        if name in self.u.dims:  # array element
            idx = self.expr()
            self.take(")")
            self.u.names.add(name)
            return "v_%s[%s]" % (name, idx)
        args = []
        while True:
            args.append(self.arg_with_shape())
            if self.peek() == ",":
                self.take()
                continue
            self.take(")")
            break
        if name in INTRINSICS:
            return "%s(%s)" % (INTRINSICS[name], ", ".join(a[0] for a in args))
        if name not in self.units or self.units[name].kind != "FUNCTION":
            raise ValueError("unknown function %s in unit %s" % (name, self.u.name))
        return "f_%s(%s)" % (name, actual_args(self.u, self.units[name], args))

    def arg_with_shape(self):
        """-> (C++ text, kind) with kind: 'var' plain variable, 'array' whole array, 'elem' array element, 'expr' anything else"""
        start = self.i
        tok = self.peek()
        if isinstance(tok, str) and re.fullmatch(r"[A-Z][A-Z0-9_]*", tok):
            nxt = self.t[self.i + 1] if self.i + 1 < len(self.t) else None
            if nxt in (",", ")", None):
                self.take()
                self.u.names.add(tok)
                return ("v_" + tok, "array" if tok in self.u.dims else "var", tok)
            if nxt == "(" and tok in self.u.dims:
                # array element, if the closing parenthesis ends the argument
                depth, j = 0, self.i + 1
                while True:
                    if self.t[j] == "(":
                        depth += 1
                    elif self.t[j] == ")":
                        depth -= 1
                        if depth == 0:
                            break
                    j += 1
                if j + 1 >= len(self.t) or self.t[j + 1] in (",", ")"):
                    txt = self.primary()
                    return (txt, "elem", tok)
        self.i = start
        return (self.expr(), "expr", None)


def actual_args(caller, callee, args):
    if len(args) != len(callee.args):
        raise ValueError("%s calls %s with %d arguments, it has %d" % (caller.name, callee.name, len(args), len(callee.args)))
    out = []
    for (txt, kind, name), dummy in zip(args, callee.args):
        dtype = callee.type_of(dummy)
        if dummy in callee.dims:  # array dummy
            if kind == "array":
                out.append(txt)
            elif kind == "elem":
                out.append("(&%s - 1)" % txt)
            else:
                raise ValueError("%s passes a scalar to array dummy %s of %s" % (caller.name, dummy, callee.name))
            if caller.type_of(name) != dtype:
                raise ValueError("type mismatch for array argument %s of %s" % (dummy, callee.name))
        elif kind in ("var", "elem"):
            if caller.type_of(name) != dtype:
                raise ValueError("%s passes %s %s to %s dummy %s of %s" % (caller.name, caller.type_of(name), name, dtype, dummy, callee.name))
            out.append(txt)
        elif kind == "array":
            raise ValueError("%s passes array %s to scalar dummy %s of %s" % (caller.name, name, dummy, callee.name))
        else:
            out.append("F_tmp<%s>(%s)" % (dtype, txt))
    return ", ".join(out)


def translate_expr(text, unit, units):
    p = ExprParser(tokenize(text), unit, units)
    e = p.expr()
    if p.peek() is not None:
        raise ValueError("trailing tokens in %r" % text)
    return e


def split_top(s, sep=","):
    out, depth, cur = [], 0, ""
    for ch in s:
        if ch == "(":
            depth += 1
        elif ch == ")":
            depth -= 1
        if ch == sep and depth == 0:
            out.append(cur)
            cur = ""
        else:
            cur += ch
    out.append(cur)
    return out


def eq_inside_parens_only(s):
    return top_level_eq(s) < 0


def top_level_eq(s):
    depth = 0
    for i, ch in enumerate(s):
        if ch == "(":
            depth += 1
        elif ch == ")":
            depth -= 1
        elif ch == "=" and depth == 0:
            return i
    return -1


HEAD = re.compile(r"^(SUBROUTINE|DOUBLEPRECISIONFUNCTION|REALFUNCTION|FUNCTION)([A-Z][A-Z0-9_]*)(?:\((.*)\))?$")
F90_TYPES = (("REAL(KIND=DP)", "double"), ("REAL(KIND=SP)", "float"), ("DOUBLEPRECISION", "double"), ("INTEGER", "int"), ("LOGICAL", "bool"),
             ("CHARACTER(LEN=*)", "const char*"), ("REAL", "float"))
DO = re.compile(r"^DO(\d+)([A-Z][A-Z0-9_]*)=(.+)$")


def declare(unit, typ, rest):
    for item in split_top(rest):
        m = re.fullmatch(r"([A-Z][A-Z0-9_]*)(?:\((.+)\))?", item)
        if not m:
            raise ValueError("declaration item %r" % item)
        if typ:
            unit.types[m.group(1)] = typ
        if m.group(2):
            unit.dims[m.group(1)] = m.group(2)


def parse_units(stmts):
    units, cur, module = {}, None, None
    for lab, s in stmts:
        if re.fullmatch(r"MODULE[A-Z][A-Z0-9_]*", s) and cur is None:
            module = cur = Unit("MODULE", s[6:], [], None)  # its specification part: variables shared by the contained units
            units["MODULE " + module.name] = module
            continue
        if s == "CONTAINS":
            cur = None
            continue
        if s.startswith("ENDMODULE"):
            module = cur = None
            continue
        if s.startswith("USE") and top_level_eq(s) < 0:
            continue
        m = HEAD.match(s)
        if m and top_level_eq(s) < 0 and (cur is None or cur.kind == "MODULE"):
            kind = "SUBROUTINE" if m.group(1) == "SUBROUTINE" else "FUNCTION"
            rtype = {"DOUBLEPRECISIONFUNCTION": "double", "REALFUNCTION": "float"}.get(m.group(1))
            cur = Unit(kind, m.group(2), [a for a in (m.group(3) or "").split(",") if a], rtype)
            if module is not None:  # host association: the module's names, unless the unit declares its own
                cur.types, cur.dims, cur.implicit = dict(module.types), dict(module.dims), dict(module.implicit)
                cur.module_vars = set(module.types) | set(module.dims)
            units[cur.name] = cur
            continue
        if cur is None:
            raise ValueError("statement outside a program unit: %r" % s)
        if s == "END" or s.startswith("ENDSUBROUTINE") or s.startswith("ENDFUNCTION"):
            cur = None
            continue
        is_assign = top_level_eq(s) >= 0 and not s.startswith("IF(") and not (DO.match(s) and len(split_top(DO.match(s).group(3))) >= 2)
        if not is_assign:
            if s.startswith("IMPLICITREAL*8"):
                cur.implicit_double = True
                continue
            if s == "IMPLICITNONE":
                continue
            m = re.fullmatch(r"IMPLICIT(REAL\(KIND=DP\)|REAL\(KIND=SP\)|INTEGER)\(([A-Z])-([A-Z])\)", s)
            if m:
                t = dict(F90_TYPES)[m.group(1)]
                for c in range(ord(m.group(2)), ord(m.group(3)) + 1):
                    cur.implicit[chr(c)] = t
                continue
            if "::" in s:
                spec, rest = s.split("::", 1)
                t = [ty for key, ty in F90_TYPES if spec == key]
                if not t:
                    raise ValueError("declaration %r" % s)
                declare(cur, t[0], rest)
                for item in split_top(rest):
                    cur.module_vars.discard(item.split("(")[0])
                continue
            if s.startswith("REAL*8"):
                declare(cur, "double", s[6:]); continue
            if s.startswith("DOUBLEPRECISION"):
                declare(cur, "double", s[15:]); continue
            if s.startswith("REAL"):
                declare(cur, "float", s[4:]); continue
            if s.startswith("LOGICAL"):
                declare(cur, "bool", s[7:]); continue
            if s.startswith("INTEGER*4"):
                declare(cur, "int", s[9:]); continue
            if s.startswith("PARAMETER("):  # named constants: evaluated once at the top of the unit, in the declared type
                for item in split_top(s[10:-1]):
                    cur.body.insert(0, ("", item))
                continue
            if s.startswith("INTEGER"):
                declare(cur, "int", s[7:]); continue
            if s.startswith("DIMENSION"):
                declare(cur, None, s[9:])
                for item in split_top(s[9:]):  # a local array hides a module variable of that name
                    if item.split("(")[0] in cur.module_vars:
                        cur.module_vars.discard(item.split("(")[0])
                        cur.types.pop(item.split("(")[0], None)
                continue
            if s.startswith("COMMON/"):
                m = re.fullmatch(r"COMMON/([A-Z0-9_]+)/(.+)", s)
                items = split_top(m.group(2))
                declare(cur, None, m.group(2))
                cur.commons.append((m.group(1), [it.split("(")[0] for it in items]))
                continue
            if s.startswith("FORMAT(") or s.startswith("WRITE(") or s.startswith("PRINT"):
                if s.startswith("FORMAT("):
                    continue
                cur.body.append((lab, "CONTINUE"))  # output statements are dropped (a label on them stays)
                continue
            if s.startswith("DATA"):
                rest = s[4:]
                while rest:
                    m = re.match(r",?([^/]+)/([^/]+)/", rest)
                    cur.data.append((split_top(m.group(1)), split_top(m.group(2))))
                    rest = rest[m.end():]
                continue
        cur.body.append((lab, s))
    return units


def translate_statement(s, unit, units):
    """one executable statement (no label handling) -> list of C++ lines; block structure through '{' / '}' lines"""
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
        cond = translate_expr(s[3:j], unit, units)
        rest = s[j + 1:]
        if rest == "THEN":
            return ["if (%s) {" % cond]
        inner = translate_statement(rest, unit, units)
        return ["if (%s) { %s }" % (cond, " ".join(inner))]
    if s.startswith("ELSEIF("):
        return ["} else if (%s) {" % translate_expr(s[7:-5], unit, units)] if s.endswith(")THEN") else None
    if s == "ELSE":
        return ["} else {"]
    if s == "ENDIF":
        return ["}"]
    if s == "CONTINUE":
        return [";"]
    if s == "RETURN":
        return ["return v_%s;" % unit.name if unit.kind == "FUNCTION" else "return;"]
    m = re.fullmatch(r"GOTO(\d+)", s)
    if m:
        return ["goto L%s;" % m.group(1)]
    eq = top_level_eq(s)
    m = DO.match(s)
    if m and len(split_top(m.group(3))) >= 2:
        lim = split_top(m.group(3))
        var = "v_" + m.group(2)
        unit.names.add(m.group(2))
        step = translate_expr(lim[2], unit, units) if len(lim) > 2 else "1"
        return ["for (%s = %s; %s <= %s; %s += %s) {" % (var, translate_expr(lim[0], unit, units), var, translate_expr(lim[1], unit, units), var, step), ("DO", m.group(1))]
    if eq >= 0:
        lhs, rhs = s[:eq], s[eq + 1:]
        m = re.fullmatch(r"([A-Z][A-Z0-9_]*)(?:\((.+)\))?", lhs)
        if not m or (m.group(2) and m.group(1) not in unit.dims):
            raise ValueError("assignment target %r in %s (statement function?)" % (lhs, unit.name))
        unit.names.add(m.group(1))
        target = "v_%s[%s]" % (m.group(1), translate_expr(m.group(2), unit, units)) if m.group(2) else "v_" + m.group(1)
        return ["%s = %s;" % (target, translate_expr(rhs, unit, units))]
    if s.startswith("OPEN(") and eq_inside_parens_only(s):
        m = re.fullmatch(r"OPEN\(UNIT=(\d+),FILE=([A-Z][A-Z0-9_]*),STATUS='OLD'\)", s)
        if not m:
            raise ValueError("OPEN form %r" % s)
        unit.names.add(m.group(2))
        return ["F_open(%s, v_%s);" % (m.group(1), m.group(2))]
    if s.startswith("CLOSE("):
        return [";"]
    m = re.fullmatch(r"READ\((\d+),\*(?:,END=(\d+))?\)(.+)", s)
    if m:  # list-directed input
        fail = "goto L%s;" % m.group(2) if m.group(2) else "F_abort(\"end of file in READ\");"
        out = ["if (!F_read_begin(%s)) { %s }" % (m.group(1), fail)]
        for item in split_top(m.group(3)):
            mm = re.fullmatch(r"\((.+),([A-Z][A-Z0-9_]*)=(.+),(.+)\)", item)
            if mm:  # implied DO
                var = "v_" + mm.group(2)
                unit.names.add(mm.group(2))
                out.append("for (%s = %s; %s <= %s; ++%s) { if (!F_read(%s, %s)) { %s } }" % (
                    var, translate_expr(mm.group(3), unit, units), var, translate_expr(mm.group(4), unit, units), var, m.group(1),
                    translate_expr(mm.group(1), unit, units), fail))
            else:
                out.append("if (!F_read(%s, %s)) { %s }" % (m.group(1), translate_expr(item, unit, units), fail))
        return out
    m = re.fullmatch(r"DO([A-Z][A-Z0-9_]*)=(.+)", s)
    if m and len(split_top(m.group(2))) >= 2:  # block DO ... END DO
        lim = split_top(m.group(2))
        var = "v_" + m.group(1)
        unit.names.add(m.group(1))
        step = translate_expr(lim[2], unit, units) if len(lim) > 2 else "1"
        return ["for (%s = %s; %s <= %s; %s += %s) {" % (var, translate_expr(lim[0], unit, units), var, translate_expr(lim[1], unit, units), var, step)]
    if s == "ENDDO":
        return ["}"]
    m = re.fullmatch(r"CALL([A-Z][A-Z0-9_]*)", s)
    if m:
        s = s + "()"
    m = re.fullmatch(r"CALL([A-Z][A-Z0-9_]*)\((.*)\)", s)
    if m:
        callee = units.get(m.group(1))
        if callee is None or callee.kind != "SUBROUTINE":
            raise ValueError("unknown subroutine %s" % m.group(1))
        p = ExprParser(tokenize(m.group(2)), unit, units)
        args = []
        while m.group(2):
            args.append(p.arg_with_shape())
            if p.peek() == ",":
                p.take()
                continue
            break
        if p.peek() is not None:
            raise ValueError("call arguments %r" % s)
        return ["f_%s(%s);" % (callee.name, actual_args(unit, callee, args))]
    raise ValueError("cannot translate %r in %s" % (s, unit.name))


PRELUDE = r"""// GENERATED by oracle/f77_to_cpp.py from the reference's Fortran at build time -- never committed.
#include <math.h>
#include <stdint.h>
#if defined(TS_SRT_MATH)
// the elementary functions of the device code (csrc/srt_math.h), for a bit-for-bit comparison with the CUDA restatement
#include "srt_math.h"
static inline double M_sin(double x) { return srt_sin(x); }
static inline double M_cos(double x) { return srt_cos(x); }
static inline double M_exp(double x) { return srt_exp(x); }
static inline double M_log(double x) { return srt_log(x); }
static inline double M_atan2(double y, double x) { return srt_atan2(y, x); }
static inline double M_atan(double x) { return srt_atan(x); }
static inline double M_tan(double x) { double s, c; srt_sincos(x, s, c); return s / c; }
static inline double M_powr(double x, double y) { return srt_pow(x, y); }
#else
static inline double M_sin(double x) { return sin(x); }
static inline double M_cos(double x) { return cos(x); }
static inline double M_exp(double x) { return exp(x); }
static inline double M_log(double x) { return log(x); }
static inline double M_atan2(double y, double x) { return atan2(y, x); }
static inline double M_atan(double x) { return atan(x); }
static inline double M_tan(double x) { return tan(x); }
static inline double M_powr(double x, double y) { return pow(x, y); }
#endif
static inline double M_sqrt(double x) { return sqrt(x); }
static inline float M_sqrt(float x) { return sqrtf(x); }
static inline double M_abs(double x) { return fabs(x); }
// REAL*4 intrinsics: libm's single-precision functions, what gfortran calls for default REAL
static inline float M_sin(float x) { return sinf(x); }
static inline float M_cos(float x) { return cosf(x); }
static inline float M_atan2(float y, float x) { return atan2f(y, x); }
static inline float M_atan(float x) { return atanf(x); }
static inline float M_asin(float x) { return asinf(x); }
static inline float M_abs(float x) { return fabsf(x); }
static inline float M_float(int i) { return (float)i; }
static inline double M_dfloat(int i) { return (double)i; }
static inline double M_mod(double a, double p) { return fmod(a, p); }
static inline float M_mod(float a, float p) { return fmodf(a, p); }
static inline int M_mod(int a, int p) { return a % p; }
static inline double M_sign(double a, double b) { return copysign(fabs(a), b); }  // SIGN(A,B): |A| with the sign bit of B (gfortran, IEEE)
static inline float M_sign(float a, float b) { return copysignf(fabsf(a), b); }
// x**n, literal n: the multiplication tree of gcc's powi table (gfc_conv_cst_int_power); table[n] for n <= 8
template <int N, class T> static inline T F_powc(T x) {
  constexpr int tbl[9] = {0, 1, 1, 2, 2, 3, 3, 4, 4};
  static_assert(N >= 1 && N <= 8, "exponent");
  if constexpr (N == 1) return x;
  else return F_powc<N - tbl[N]>(x) * F_powc<tbl[N]>(x);
}
// x**i, integer variable: libgcc __powidf2
static inline double F_pow(double x, int m) {
  unsigned n = m < 0 ? -(unsigned)m : (unsigned)m;
  double y = (n % 2) ? x : 1;
  while (n >>= 1) { x = x * x; if (n % 2) y *= x; }
  return m < 0 ? 1 / y : y;
}
static inline double F_pow(double x, double y) { return M_powr(x, y); }
static inline double F_pow(double x, float y) { return M_powr(x, (double)y); }
template <class T> static inline T& F_tmp(T&& v) { return v; }
#define TL static thread_local
// list-directed sequential input (READ (u,*) ...): every READ starts a new record, items are separated by blanks or commas
// and continue onto following records; `d` exponents accepted
#include <stdio.h>
#include <stdlib.h>
#include <string>
struct F_Unit { FILE* f = nullptr; std::string line; size_t pos = 0; bool failed = false; };
static F_Unit F_units[16];
static void F_abort(const char* what) { fprintf(stderr, "translated Fortran: %s\n", what); abort(); }
static void F_open(int u, const char* name) { if (F_units[u].f) fclose(F_units[u].f); F_units[u] = F_Unit(); F_units[u].f = fopen(name, "r"); F_units[u].failed = !F_units[u].f; }
static bool F_next_line(F_Unit& U) {
  if (!U.f) return false;
  U.line.clear(); U.pos = 0;
  int c; bool any = false;
  while ((c = fgetc(U.f)) != EOF) { any = true; if (c == '\n') break; if (c != '\r') U.line.push_back((char)c); }
  return any;
}
static bool F_read_begin(int u) { return F_next_line(F_units[u]); }
template <class T> static bool F_read(int u, T& v) {
  F_Unit& U = F_units[u];
  for (;;) {
    while (U.pos < U.line.size() && (U.line[U.pos] == ' ' || U.line[U.pos] == ',' || U.line[U.pos] == '\t')) ++U.pos;
    if (U.pos < U.line.size()) break;
    if (!F_next_line(U)) return false;
  }
  std::string tok;
  while (U.pos < U.line.size() && U.line[U.pos] != ' ' && U.line[U.pos] != ',' && U.line[U.pos] != '\t') tok.push_back(U.line[U.pos++]);
  for (char& ch : tok) if (ch == 'd' || ch == 'D') ch = 'e';
  v = (T)strtod(tok.c_str(), nullptr);
  return true;
}
"""


def data_values(vals):
    out = []
    for v in vals:
        m = re.fullmatch(r"(\d+)\*(.+)", v)
        rep, lit = (int(m.group(1)), m.group(2)) if m else (1, v)
        sign = ""
        if lit[0] in "+-":
            sign, lit = lit[0], lit[1:]
        out += [sign + number_to_cpp(lit)] * rep
    return out


def emit_unit(u, units, common_types):
    rt = "void" if u.kind == "SUBROUTINE" else (u.rtype or u.type_of(u.name))
    params = ", ".join(("%s v_%s" if u.type_of(a) == "const char*" else "%s* v_%s" if a in u.dims else "%s& v_%s") % (u.type_of(a), a) for a in u.args)
    body_lines, do_stack = [], []
    for lab, s in u.body:
        if lab:
            body_lines.append("L%s: ;" % lab)
        for ln in translate_statement(s, u, units):
            if isinstance(ln, tuple):
                do_stack.append(ln[1])
            else:
                body_lines.append(ln)
        while lab and do_stack and do_stack[-1] == lab:
            do_stack.pop()
            body_lines.append("}")
    if do_stack:
        raise ValueError("unterminated DO in %s" % u.name)
    in_common = {}
    size_of = {"int": 4, "float": 4, "bool": 4, "double": 8}
    for block, names in u.commons:
        off = 0
        for n in names:
            sz = size_of[u.type_of(n)]
            off = (off + sz - 1) // sz * sz
            in_common[n] = (block, off)
            off += sz * (int(u.dims[n]) if n in u.dims else 1)
        common_types[block] = max(common_types.get(block, 0), off)
    names = set(u.names) | set(u.dims) | {n for ns, _ in u.data for n in ns}
    if u.kind == "FUNCTION":
        names.add(u.name)
    init = {}
    once = []
    for ns, vals in u.data:
        vals = data_values(vals)
        k = 0
        for n in ns:
            cnt = int(u.dims[n]) if n in u.dims else 1
            chunk = vals[k:k + cnt]
            k += cnt
            if len(chunk) != cnt:
                raise ValueError("DATA for %s in %s: too few values" % (n, u.name))
            if n in in_common or n in u.args:
                once += ["v_%s = %s;" % (n, chunk[0])] if cnt == 1 else ["v_%s[%d] = %s;" % (n, i + 1, c) for i, c in enumerate(chunk)]
            else:
                init[n] = chunk
        if k != len(vals):
            raise ValueError("DATA in %s: %d values left over" % (u.name, len(vals) - k))
    decl = []
    for n in sorted(names):
        if n in u.args or n in u.module_vars:
            continue
        t = u.type_of(n)
        if n in in_common and n in u.dims:
            decl.append("%s* v_%s = (%s*)(C_%s + %d) - 1;" % (t, n, t, in_common[n][0], in_common[n][1]))
        elif n in in_common:
            decl.append("%s& v_%s = *(%s*)(C_%s + %d);" % (t, n, t, in_common[n][0], in_common[n][1]))
        elif n in u.dims:
            size = int(u.dims[n])
            decl.append("TL %s v_%s[%d] = {0%s};" % (t, n, size + 1, "".join(", " + v for v in init.get(n, []))))
        elif n in units and n != u.name:
            continue
        else:
            decl.append("TL %s v_%s = %s;" % (t, n, init[n][0] if n in init else "0"))
    if once:
        decl.append("TL bool data_done = false; if (!data_done) { data_done = true; %s }" % " ".join(once))
    tail = ["return v_%s;" % u.name] if u.kind == "FUNCTION" else []
    hoist = ""
    if u.name == "T04_S":  # its coefficient array is needed by the EXTERN wrapper
        keep = [d for d in decl if d.startswith("TL double v_A[")]
        decl = [d for d in decl if d not in keep]
        hoist = keep[0] + "\ndouble* ts05_ref_A() { return v_A; }\n"
    text = hoist + "%s f_%s(%s) {\n  %s\n  %s\n}\n" % (rt, u.name, params, "\n  ".join(decl), "\n  ".join(body_lines + tail))
    proto = "%s f_%s(%s);" % (rt, u.name, params)
    return proto, text


TS05_WRAPPERS = """
// C entry points for the tests (ctypes): the model as the adapters call it, and its modules one by one
extern "C" void ts05_ref_t04_s(const float* parmod, float ps, float x, float y, float z, float* b) {
  int iopt = 0; float p[11]; for (int i = 0; i < 10; ++i) p[i + 1] = parmod[i];
  float bx = 0, by = 0, bz = 0;
  f_T04_S(iopt, p, ps, x, y, z, bx, by, bz);
  b[0] = bx; b[1] = by; b[2] = bz;
}
// EXTERN in double precision: par = pdyn, dst*, byimf, bzimf, w1..w6 (dst* as T04_s forms it); out[30] = total field, then
// the module fields CF, T1, T2, SRC, PRC, R11, R12, R21, R22 before their amplitudes
extern "C" void ts05_ref_extern(const double* par, double ps, double x, double y, double z, double* out) {
  int z0 = 0, z1 = 0, z2 = 0, z3 = 0, ntot = 69;
  TL double bximf = 0;
  double p[10]; for (int i = 0; i < 10; ++i) p[i] = par[i];
  double o[33] = {0};
  // the coefficient array A lives in T04_s: fetch it through one call of the model
  { int iopt = 0; float pm[11] = {0, 2.f, -20.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, fx = -5.f, fy = 0.f, fz = 0.f, fps = 0.f, b0, b1, b2; f_T04_S(iopt, pm, fps, fx, fy, fz, b0, b1, b2); }
  f_EXTERN(z0, z1, z2, z3, ts05_ref_A(), ntot, p[0], p[1], bximf, p[2], p[3], p[4], p[5], p[6], p[7], p[8], p[9], ps, x, y, z,
           o[3], o[4], o[5], o[6], o[7], o[8], o[9], o[10], o[11], o[12], o[13], o[14], o[15], o[16], o[17], o[18], o[19], o[20],
           o[21], o[22], o[23], o[24], o[25], o[26], o[27], o[28], o[29], o[30], o[31], o[32], o[0], o[1], o[2]);
  for (int i = 0; i < 30; ++i) out[i] = o[i];
}
"""

GEOPACK_WRAPPERS = """
// C entry points for the tests (ctypes): RECALC_08 as TSY_RECALC calls it (geopack0508_adapter.for:21-30), then the state of
// COMMON /GEOPACK1/ (34 words) and /GEOPACK2/ (G, H, REC: 3 x 105), and IGRF_GSW_08 (the adapters' IGRF_GSM)
extern "C" void geopack_ref_recalc(int iyear, int iday, int ihour, int min, int isec, float* geopack1, float* geopack2) {
  float vx = -400.0f, vy = 0.0f, vz = 0.0f;
  f_RECALC_08(iyear, iday, ihour, min, isec, vx, vy, vz);
  for (int i = 0; i < 34; ++i) geopack1[i] = ((float*)C_GEOPACK1)[i];
  for (int i = 0; i < 315; ++i) geopack2[i] = ((float*)C_GEOPACK2)[i];
}
extern "C" void geopack_ref_igrf_gsw(float x, float y, float z, float* b) {
  float bx = 0, by = 0, bz = 0;
  f_IGRF_GSW_08(x, y, z, bx, by, bz);
  b[0] = bx; b[1] = by; b[2] = bz;
}
"""


NGO_WRAPPERS = """
// C entry points for the tests (ctypes): readinput, and dens the way funcPlasmaParams drives it (ngo_dens_model_adapter.f95:
// 120-133: z(1), z(2), latitu set, `call dens`, ani(1:4) read)
extern "C" int ngo_ref_readinput(const char* filename) { f_READINPUT(filename); return F_units[4].failed ? -1 : 0; }
extern "C" void ngo_ref_dens(double z1, double z2, double latitu, double* ani) {
  v_Z[1] = z1; v_Z[2] = z2; v_LATITU = latitu;
  f_DENS();
  for (int i = 0; i < 4; ++i) ani[i] = v_ANI[i + 1];
}
extern "C" void ngo_ref_state(double* out) {
  const double s[12] = {v_ANE0, v_RBASE, v_R0, (double)v_NUM, (double)v_KDUCTS, v_THERM, v_RZERO, v_SCBOT, v_LK, v_EXPK, v_DDK, v_PI};
  for (int i = 0; i < 12; ++i) out[i] = s[i];
}
"""


XFORM_WRAPPERS = """
// C entry points for the tests (ctypes): the two transformations the adapters call (ngo_dens_model_adapter.f95:145,202)
extern "C" void xform_ref_sm_to_gsm(int yearday, int msec, const double* x, double* out) {
  int itime[3] = {0, yearday, msec}; double xi[4] = {0, x[0], x[1], x[2]}, xo[4] = {0, 0, 0, 0};
  f_SM_TO_GSM_D(itime, xi, xo);
  out[0] = xo[1]; out[1] = xo[2]; out[2] = xo[3];
}
extern "C" void xform_ref_gsm_to_sm(int yearday, int msec, const double* x, double* out) {
  int itime[3] = {0, yearday, msec}; double xi[4] = {0, x[0], x[1], x[2]}, xo[4] = {0, 0, 0, 0};
  f_GSM_TO_SM_D(itime, xi, xo);
  out[0] = xo[1]; out[1] = xo[2]; out[2] = xo[3];
}
"""


XFORMD_WRAPPERS = """
// the Fortran entry points of libxformd the reference's damping code links (damping/include/coord_transforms.h:31-48), under the
// names gfortran gives them; arguments by reference, arrays 1-based inside the translation
extern "C" void sm_to_mag_d_(int* itime, double* x_in, double* x_out) { f_SM_TO_MAG_D(itime - 1, x_in - 1, x_out - 1); }
extern "C" void geo_to_mag_d_(int* itime, double* x_in, double* x_out) { f_GEO_TO_MAG_D(itime - 1, x_in - 1, x_out - 1); }
extern "C" void mag_to_geo_d_(int* itime, double* x_in, double* x_out) { f_MAG_TO_GEO_D(itime - 1, x_in - 1, x_out - 1); }
extern "C" void cart_to_pol_d_(double* x_in, double* lat, double* lon, double* radial) { f_CART_TO_POL_D(x_in - 1, *lat, *lon, *radial); }
extern "C" void pol_to_cart_d_(double* lat, double* lon, double* radial, double* x_out) { f_POL_TO_CART_D(*lat, *lon, *radial, x_out - 1); }
"""


def main(src, dst, only=None, wrappers="ts05"):
    free = wrappers in ("ngo", "xform", "xformd")
    stmts = []
    for one in src.split(","):
        stmts += read_statements_free(one) if free else read_statements(one)
    units = parse_units(stmts)
    if only:
        units = {k: v for k, v in units.items() if k in only}
    common_types, protos, texts = {}, [], []
    for u in units.values():
        if u.kind == "MODULE":  # module variables: one object each, shared by the contained units
            for n in sorted(set(u.types) | set(u.dims)):
                protos.append("TL %s v_%s%s;" % (u.type_of(n), n, "[%d] = {0}" % (int(u.dims[n]) + 1) if n in u.dims else " = 0"))
            continue
        p, t = emit_unit(u, units, common_types)
        protos.append(p)
        texts.append(t)
    commons = []
    for block, size in sorted(common_types.items()):  # storage sequences: members are matched by byte offset, as in Fortran
        commons.append("alignas(8) TL unsigned char C_%s[%d] = {0};" % (block, (size + 7) // 8 * 8))
    with open(dst, "w") as f:
        f.write(PRELUDE + "\n" + "\n".join(commons) + ("\ndouble* ts05_ref_A();\n\n" if wrappers == "ts05" else "\n\n") + "\n".join(protos) + "\n\n" + "\n".join(texts))
        f.write({"ts05": TS05_WRAPPERS, "geopack": GEOPACK_WRAPPERS, "ngo": NGO_WRAPPERS, "xform": XFORM_WRAPPERS, "xformd": XFORMD_WRAPPERS}[wrappers])


if __name__ == "__main__":
    opts = dict(a[2:].split("=", 1) for a in sys.argv[3:] if a.startswith("--"))
    main(sys.argv[1], sys.argv[2], only=set(opts["units"].split(",")) if "units" in opts else None, wrappers=opts.get("wrappers", "ts05"))
