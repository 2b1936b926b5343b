#!/usr/bin/env python
"""Translator of the FORTRAN 77 dialect LAPACK 3.2.1 is written in (explicit typing, COMPLEX*16, CHARACTER*1 options,
assumed-size and two-dimensional dummy arrays, labelled DO / GO TO) to C++.

TEST INFRASTRUCTURE (oracle/): the reference's `csvd` (fortran/blas.f95:180-221) calls `zgesvd` of the LAPACK 3.2.1 it vendors
under lapack-3.2.1/ and links (fortran/Makefile:48).  No Fortran compiler exists here, so the routines under that call
(SRC/*.f, BLAS/SRC/*.f) are translated statement by statement where they lie and compiled into oracle/_ref/liblapack_ref.so
-- the translated text never enters the repository.  Every operation, its order and its type are kept:
  * scalars are passed by reference, arrays as pointers with 1-based indexing, A(I,J) = a[I + (J-1)*LDA];
  * COMPLEX*16 is std::complex<double> compiled with -fcx-fortran-rules; REAL op COMPLEX is the operators of f90rt.h;
  * a DO loop has its trip count fixed on entry and leaves its variable one step past the last value;
  * x**2 is x*x, x**i with an integer variable is libgcc's __powidf2.
What is NOT translated is stated in the prelude: LSAME, XERBLA, ILAENV (block sizes: no result depends on them for a
3 x 3 matrix, every blocked branch falls back to the unblocked routine) and DLAMCH (the IEEE double constants its
run-time probing finds); routines of branches a 3 x 3 'A','A' call never enters abort if entered.
Usage: lapack_to_cpp.py out.cpp file.f [file.f ...]
"""
import os
import re
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from f77_to_cpp import read_statements, split_top, top_level_eq  # noqa: E402

TYPES = (("DOUBLEPRECISION", "double"), ("DOUBLECOMPLEX", "cd"), ("COMPLEX*16", "cd"), ("INTEGER", "int"), ("LOGICAL", "bool"), ("CHARACTER*1", "const char*"),
         ("CHARACTER*(*)", "const char*"), ("CHARACTER", "const char*"))
INTRINSICS = {"ABS": "F_abs", "MAX": "F_max", "MIN": "F_min", "SQRT": "F_sqrt", "SIGN": "F_sign", "DBLE": "F_dble", "DCMPLX": "F_dcmplx",
              "DCONJG": "F_conjg", "DIMAG": "F_dimag", "INT": "F_int", "LOG": "F_log", "MOD": "F_mod"}
STUBS = {"LSAME", "XERBLA", "ILAENV", "DLAMCH"}  # hand-written in the prelude
LOGOPS = {".LT.": "<", ".GT.": ">", ".LE.": "<=", ".GE.": ">=", ".EQ.": "==", ".NE.": "!="}
HEAD = re.compile(r"^(?:(DOUBLEPRECISION|DOUBLECOMPLEX|COMPLEX\*16|INTEGER|LOGICAL))?(SUBROUTINE|FUNCTION)([A-Z][A-Z0-9_]*)(?:\((.*)\))?$")


def tokenize(s):
    toks, i = [], 0
    while i < len(s):
        c = s[i]
        if c == "'":
            j = s.index("'", i + 1)
            toks.append(("STR", s[i + 1:j]))
            i = j + 1
            continue
        m = re.match(r"\.(LT|GT|LE|GE|EQ|NE|AND|OR|NOT|TRUE|FALSE)\.", s[i:])
        if m:
            toks.append(m.group(0))
            i += m.end()
            continue
        if c.isalpha():
            m = re.match(r"[A-Z][A-Z0-9_]*", s[i:])
            toks.append(m.group(0))
            i += m.end()
            continue
        if c.isdigit() or (c == "." and s[i + 1:i + 2].isdigit()):
            j = i
            while j < len(s) and s[j].isdigit():
                j += 1
            if j < len(s) and s[j] == "." and not re.match(r"\.(LT|GT|LE|GE|EQ|NE|AND|OR|NOT)\.", s[j:]):
                j += 1
                while j < len(s) and s[j].isdigit():
                    j += 1
            m = re.match(r"[DE][+-]?\d+", s[j:])
            if m:
                j += m.end()
            toks.append(("NUM", s[i:j]))
            i = j
            continue
        if s[i:i + 2] in ("**", "//"):
            toks.append(s[i:i + 2])
            i += 2
            continue
        if c in "(),+-*/=:":
            toks.append(c)
            i += 1
            continue
        raise ValueError("cannot tokenize %r at %r" % (s, s[i:]))
    return toks


def number(txt):
    if re.fullmatch(r"\d+", txt):
        return txt
    t = txt.replace("D", "e").replace("E", "e")
    mant = t.split("e")[0]
    if "." not in mant:
        t = t.replace(mant, mant + ".", 1)
    if "D" not in txt:
        raise ValueError("single precision literal %s" % txt)  # none is expected in the double precision routines
    return t


class Unit:
    def __init__(self, kind, name, args, rtype):
        self.kind, self.name, self.args, self.rtype = kind, name, args, rtype
        self.types, self.dims, self.params, self.body = {}, {}, {}, []

    def type_of(self, name):
        if name == self.name and self.rtype:
            return self.rtype
        if name not in self.types:
            raise ValueError("%s: %s has no declared type" % (self.name, name))
        return self.types[name]


class Expr:
    def __init__(self, toks, unit, units):
        self.t, self.i, self.u, self.units = toks, 0, unit, units

    def peek(self, k=0):
        return self.t[self.i + k] if self.i + k < len(self.t) else None

    def take(self, want=None):
        tok = self.peek()
        if want is not None and tok != want:
            raise ValueError("%s: expected %r, got %r in %r" % (self.u.name, want, tok, self.t))
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
        if self.peek() in LOGOPS:
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
            if isinstance(tok, tuple) and tok[0] == "NUM" and tok[1] == "2":
                self.take()
                return "F_sq(%s)" % base
            return "F_pow(%s, %s)" % (base, self.pow_())
        return base

    def index(self, name):
        """after '(' of an array reference: -> C++ index text (1-based storage)"""
        subs = []
        while True:
            subs.append(self.expr())
            if self.peek() == ",":
                self.take()
                continue
            self.take(")")
            break
        dims = self.u.dims[name]
        if len(subs) != len(dims):
            raise ValueError("%s: %s has %d dimensions, used with %d" % (self.u.name, name, len(dims), len(subs)))
        if len(subs) == 1:
            return subs[0]
        return "(%s) + ((%s) - 1) * (long)(%s)" % (subs[0], subs[1], translate(dims[0], self.u, self.units))

    def primary(self):
        tok = self.take()
        if tok == "(":
            e = self.expr()
            if self.peek() == ",":  # a COMPLEX constant (re, im)
                self.take()
                im = self.expr()
                self.take(")")
                return "cd(%s, %s)" % (e, im)
            self.take(")")
            return "(%s)" % e
        if isinstance(tok, tuple):
            return number(tok[1]) if tok[0] == "NUM" else '"%s"' % tok[1]
        if tok in (".TRUE.", ".FALSE."):
            return "true" if tok == ".TRUE." else "false"
        name = tok
        if self.peek() != "(":
            return "v_" + name
        self.take("(")
        if name in self.u.dims:
            return "v_%s[%s]" % (name, self.index(name))
        if name in INTRINSICS:
            args = []
            while True:
                args.append(self.expr())
                if self.peek() == ",":
                    self.take()
                    continue
                self.take(")")
                break
            return "%s(%s)" % (INTRINSICS[name], ", ".join(args))
        return "f_%s(%s)" % (name, self.call_args(name))

    def call_args(self, callee_name):
        """actual arguments after '(' up to the closing ')'"""
        callee = self.units.get(callee_name)
        args = []
        if self.peek() == ")":
            self.take()
        else:
            while True:
                args.append(self.one_arg())
                if self.peek() == ",":
                    self.take()
                    continue
                self.take(")")
                break
        if callee is None:
            if callee_name not in STUBS and callee_name not in ABSENT:
                raise ValueError("%s calls %s, which is neither translated nor declared absent" % (self.u.name, callee_name))
            return ", ".join("F_tmp(%s)" % txt if kind in ("expr", "str") else txt for txt, kind, name in args)
        if len(args) != len(callee.args):
            raise ValueError("%s calls %s with %d arguments, it has %d" % (self.u.name, callee_name, len(args), len(callee.args)))
        out = []
        for (txt, kind, name), dummy in zip(args, callee.args):
            dtype = callee.type_of(dummy)
            if dummy in callee.dims:
                if kind == "expr" or kind == "str":
                    raise ValueError("%s passes an expression to array dummy %s of %s" % (self.u.name, dummy, callee_name))
                ptr = txt if kind == "array" else "(&%s - 1)" % txt
                if self.u.type_of(name) != dtype:
                    raise ValueError("%s passes %s %s to %s array dummy %s of %s" % (self.u.name, self.u.type_of(name), name, dtype, dummy, callee_name))
                out.append(ptr)
            elif kind in ("var", "elem"):
                if self.u.type_of(name) != dtype:
                    raise ValueError("%s passes %s %s to %s dummy %s of %s" % (self.u.name, self.u.type_of(name), name, dtype, dummy, callee_name))
                out.append(txt)
            elif kind == "array":  # an array name for a scalar dummy: its first element
                out.append("%s[1]" % txt)
            else:
                out.append("F_tmp<%s>((%s)(%s))" % (dtype, dtype, txt))
        return ", ".join(out)

    def one_arg(self):
        """-> (text, kind, name): 'var' scalar variable, 'array' whole array, 'elem' array element, 'str' literal, 'expr' the rest"""
        start, tok = self.i, self.peek()
        if isinstance(tok, tuple) and tok[0] == "STR" and self.peek(1) in (",", ")"):
            self.take()
            return ('"%s"' % tok[1], "str", None)
        if isinstance(tok, str) and re.fullmatch(r"[A-Z][A-Z0-9_]*", tok):
            nxt = self.peek(1)
            if nxt in (",", ")") and tok not in self.u.params:
                self.take()
                return ("v_" + tok, "array" if tok in self.u.dims else "var", tok)
            if nxt == "(" and tok in self.u.dims:
                depth, j = 0, self.i + 1
                while True:
                    if self.t[j] == "(":
                        depth += 1
                    elif self.t[j] == ")":
                        depth -= 1
                        if depth == 0:
                            break
                    j += 1
                if self.t[j + 1] in (",", ")"):
                    return (self.primary(), "elem", tok)
        self.i = start
        # CHARACTER concatenation appears only in the option string handed to ILAENV, which no result depends on
        if "//" in self.t[start:]:
            depth, j = 0, start
            while not (depth == 0 and self.t[j] in (",", ")")):
                depth += {"(": 1, ")": -1}.get(self.t[j], 0) if isinstance(self.t[j], str) else 0
                j += 1
            if "//" in self.t[start:j]:
                self.i = j
                return ('" "', "str", None)
        return (self.expr(), "expr", None)


ABSENT = set()


def translate(text, unit, units):
    p = Expr(tokenize(text), unit, units)
    e = p.expr()
    if p.peek() is not None:
        raise ValueError("%s: trailing tokens in %r" % (unit.name, text))
    return e


def parse(stmts):
    units, cur = {}, None
    for lab, s in stmts:
        if cur is None:
            m = HEAD.match(s)
            if not m:
                raise ValueError("unit header expected, got %r" % s)
            rtype = dict(TYPES).get(m.group(1)) if m.group(1) else None
            cur = Unit(m.group(2), m.group(3), [a for a in (m.group(4) or "").split(",") if a], rtype)
            continue
        if s in ("END", "ENDFUNCTION", "ENDSUBROUTINE") and not lab:
            units[cur.name] = cur
            cur = None
            continue
        if s.startswith(("EXTERNAL", "INTRINSIC", "IMPLICITNONE")) and top_level_eq(s) < 0:
            continue
        if s.startswith("PARAMETER(") and s.endswith(")"):
            for item in split_top(s[10:-1]):
                k = top_level_eq(item)
                cur.params[item[:k]] = item[k + 1:]
            continue
        typ = next(((k, v) for k, v in TYPES if s.startswith(k)), None)
        if typ and top_level_eq(s) < 0 and not cur.body:
            for item in split_top(s[len(typ[0]):]):
                m = re.fullmatch(r"([A-Z][A-Z0-9_]*)(?:\((.+)\))?", item)
                cur.types[m.group(1)] = typ[1]
                if m.group(2):
                    cur.dims[m.group(1)] = split_top(m.group(2))
            continue
        cur.body.append((lab, s))
    return units


def statement(s, u, units, state):
    """-> list of C++ lines; state: {'do': [(label or None, end text)], 'n': counter}"""
    tr = lambda x: translate(x, u, units)  # noqa: E731
    if s.startswith("IF("):
        depth, j = 0, 2
        while True:
            depth += {"(": 1, ")": -1}.get(s[j], 0)
            if depth == 0:
                break
            j += 1
        cond, rest = tr(s[3:j]), s[j + 1:]
        if rest == "THEN":
            return ["if (%s) {" % cond]
        return ["if (%s) { %s }" % (cond, " ".join(statement(rest, u, units, state)))]
    if s.startswith("ELSEIF(") and s.endswith(")THEN"):
        return ["} else if (%s) {" % tr(s[7:-5])]
    if s == "ELSE":
        return ["} else {"]
    if s == "ENDIF":
        return ["}"]
    m = re.fullmatch(r"DO(\d*)([A-Z][A-Z0-9_]*)=(.+)", s)
    if m and len(split_top(m.group(3))) >= 2 and (m.group(2) in u.types or m.group(2) == u.name):
        lim = split_top(m.group(3))
        state["n"] += 1
        k, v = state["n"], "v_" + m.group(2)
        step = tr(lim[2]) if len(lim) > 2 else "1"
        state["do"].append((m.group(1) or None, "}}"))
        return ["{ const int F_lo%d = %s, F_hi%d = %s, F_st%d = %s; long F_n%d = F_trips(F_lo%d, F_hi%d, F_st%d);" % (k, tr(lim[0]), k, tr(lim[1]), k, step, k, k, k, k),
                "for (%s = F_lo%d; F_n%d > 0; --F_n%d, %s += F_st%d) {" % (v, k, k, k, v, k)]
    m = re.fullmatch(r"DOWHILE\((.+)\)", s)
    if m:
        state["do"].append((None, "}"))
        return ["while (%s) {" % tr(m.group(1))]
    if s in ("ENDDO",):
        lab, end = state["do"].pop()
        return [end]
    if s == "EXIT":
        return ["break;"]
    if s == "CYCLE":
        return ["continue;"]
    if s == "CONTINUE":
        return [";"]
    m = re.fullmatch(r"GOTO(\d+)", s)
    if m:
        return ["goto L%s;" % m.group(1)]
    if s == "RETURN":
        return ["return%s;" % (" v_" + u.name if u.kind == "FUNCTION" else "")]
    m = re.fullmatch(r"CALL([A-Z][A-Z0-9_]*)(?:\((.*)\))?", s)
    if m:
        p = Expr(tokenize("(" + (m.group(2) or "") + ")"), u, units)
        p.take("(")
        return ["f_%s(%s);" % (m.group(1), p.call_args(m.group(1)))]
    k = top_level_eq(s)
    if k >= 0:
        lhs_type = u.type_of(re.match(r"[A-Z][A-Z0-9_]*", s).group(0))
        conv = {"double": "F_to_double", "int": "F_to_int"}.get(lhs_type)  # COMPLEX assigned to REAL keeps the real part, REAL to INTEGER truncates
        return ["%s = %s;" % (tr(s[:k]), "%s(%s)" % (conv, tr(s[k + 1:])) if conv else tr(s[k + 1:]))]
    raise ValueError("%s: cannot translate %r" % (u.name, s))


def emit(u, units):
    def ctype(n):
        return u.type_of(n)
    params = ", ".join(("%s* v_%s" if a in u.dims else "%s& v_%s") % (ctype(a), a) for a in u.args)
    rt = "void" if u.kind == "SUBROUTINE" else u.rtype
    decl = []
    for n, expr in u.params.items():
        decl.append("const %s v_%s = %s;" % (ctype(n), n, translate(expr, u, units)))
    for n, t in u.types.items():
        if n in u.args or n in u.params or n == u.name or n in units or n in STUBS or n in ABSENT:
            continue
        if n in u.dims:
            size = " * ".join("(%s)" % translate(d, u, units) for d in u.dims[n])
            decl.append("%s a_%s[(%s) + 1] = {}; %s* v_%s = a_%s;" % (t, n, size, t, n, n))
        else:
            decl.append("%s v_%s = %s;" % (t, n, {"bool": "false", "const char*": '""'}.get(t, "0")))
    if u.kind == "FUNCTION":
        decl.append("%s v_%s = %s;" % (rt, u.name, "false" if rt == "bool" else "0"))
    body, state = [], {"do": [], "n": 0}
    for lab, s in u.body:
        closes = 0
        if lab:
            body.append("L%s:;" % lab)
        body += statement(s, u, units, state)
        while lab and state["do"] and state["do"][-1][0] == lab:  # the terminal statement of (possibly several) labelled DO loops
            body.append(state["do"].pop()[1])
    if state["do"]:
        raise ValueError("%s: unterminated DO" % u.name)
    tail = ["return v_%s;" % u.name] if u.kind == "FUNCTION" else []
    proto = "%s f_%s(%s)" % (rt, u.name, params)
    return proto + ";", "%s {\n  %s\n  %s\n}\n" % (proto, "\n  ".join(decl), "\n  ".join(body + tail))


PRELUDE = r"""// GENERATED by oracle/lapack_to_cpp.py from the reference's lapack-3.2.1 at build time -- never committed.
#include <float.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "f90rt.h"   // cd and the Fortran-rule REAL op COMPLEX operators

// (F_tmp, F_pow(double, int) come from f90rt.h)
inline long F_trips(int lo, int hi, int st) { long n = ((long)hi - lo + st) / st; return n > 0 ? n : 0; }
inline double F_to_double(double x) { return x; }
inline double F_to_double(const cd& z) { return z.real(); }
inline int F_to_int(int x) { return x; }
inline int F_to_int(double x) { return (int)x; }
inline double F_sq(double x) { return x * x; }
inline int F_sq(int x) { return x * x; }
inline cd F_sq(const cd& x) { return x * x; }
inline double F_abs(double x) { return fabs(x); }
inline int F_abs(int x) { return x < 0 ? -x : x; }
inline double F_abs(const cd& z) { return std::abs(z); }       // ABS of a COMPLEX*16: libm's cabs, as gfortran calls it
inline double F_sqrt(double x) { return sqrt(x); }
inline double F_log(double x) { return log(x); }
inline double F_dble(double x) { return x; }
inline double F_dble(int x) { return (double)x; }
inline double F_dble(const cd& z) { return z.real(); }
inline double F_dimag(const cd& z) { return z.imag(); }
inline cd F_conjg(const cd& z) { return cd(z.real(), -z.imag()); }
inline cd F_dcmplx(double re) { return cd(re, 0.0); }
inline cd F_dcmplx(int re) { return cd((double)re, 0.0); }
inline cd F_dcmplx(const cd& z) { return z; }
inline cd F_dcmplx(double re, double im) { return cd(re, im); }
inline int F_int(double x) { return (int)x; }
inline int F_mod(int a, int b) { return a % b; }
inline double F_sign(double a, double b) { return copysign(fabs(a), b); }  // SIGN(A,B) with gfortran: the sign bit of B, also of -0.0
inline int F_sign(int a, int b) { int m = a < 0 ? -a : a; return b < 0 ? -m : m; }
// MAX / MIN as gfortran expands them for non-NaN operands (first operand kept on ties)
inline double F_max(double a, double b) { return b > a ? b : a; }
inline int F_max(int a, int b) { return b > a ? b : a; }
template <class T, class... R> T F_max(T a, T b, R... r) { return F_max(F_max(a, b), r...); }
inline double F_min(double a, double b) { return b < a ? b : a; }
inline int F_min(int a, int b) { return b < a ? b : a; }
template <class T, class... R> T F_min(T a, T b, R... r) { return F_min(F_min(a, b), r...); }
inline int F_pow(int x, int m) { int y = 1; for (int i = 0; i < m; ++i) y *= x; return y; }

// ---- not translated (stated in lapack_to_cpp.py)
template <class A, class B> bool f_LSAME(A&& a, B&& b) {   // BLAS/SRC/lsame.f: case-insensitive comparison of the first characters
  int x = (unsigned char)a[0], y = (unsigned char)b[0];
  if (x >= 'a' && x <= 'z') x -= 32;
  if (y >= 'a' && y <= 'z') y -= 32;
  return x == y;
}
template <class A, class B> void f_XERBLA(A&& name, B&& info) { fprintf(stderr, "lapack_ref: XERBLA %s %d\n", (const char*)name, (int)info); abort(); }
// ILAENV (SRC/ilaenv.f): ISPEC = 1 block size 32 for the routines met here (64 for xGEMM-rich ones does not occur), 2 minimum
// block size 2, 3 crossover 128, 6 INT(MIN(N1,N2)*1.6).  For a 3 x 3 matrix every consumer falls to its unblocked branch.
template <class S, class O> int f_ILAENV(int& ispec, S&& name, O&& opts, int& n1, int& n2, int& n3, int& n4) {
  (void)name; (void)opts; (void)n3; (void)n4;
  switch (ispec) {
    case 1: return 32;
    case 2: return 2;
    case 3: return 128;
    case 6: return (int)((float)(n1 < n2 ? n1 : n2) * 1.6E0f);
    default: fprintf(stderr, "lapack_ref: ILAENV ispec %d\n", ispec); abort();
  }
}
// DLAMCH (INSTALL/dlamch.f probes the arithmetic at run time): what it finds for IEEE double with round-to-nearest
template <class C> double f_DLAMCH(C&& cmach) {
  const double eps = DBL_EPSILON * 0.5, sfmin = DBL_MIN, base = 2.0;   // rnd = 1: eps = base**(1-t)/2; 1/huge < tiny so sfmin = tiny
  switch (cmach[0] & ~32) {
    case 'E': return eps;
    case 'S': return sfmin;
    case 'B': return base;
    case 'P': return eps * base;
    case 'N': return 53.0;
    case 'R': return 1.0;
    case 'M': return -1021.0;
    case 'U': return DBL_MIN;
    case 'L': return 1024.0;
    case 'O': return DBL_MAX;
    default: fprintf(stderr, "lapack_ref: DLAMCH %c\n", cmach[0]); abort();
  }
}
"""

ENTRY = r"""
extern "C" {
// csvd of the reference (fortran/blas.f95:180-221) on a 3 x 3 matrix: workspace query, then zgesvd('A', 'A', ...).
// a, u, vt: column-major (re, im) pairs; returns INFO
int lapack_ref_csvd3(const double* a_re_im, double* s, double* u_re_im, double* vt_re_im) {
  cd A[10], U[10], VT[10], worktmp[2], work[4096];
  double S[4], rwork[16];
  int m = 3, n = 3, lda = 3, info = 0, lwork = -1;
  for (int k = 0; k < 9; ++k) A[k + 1] = cd(a_re_im[2 * k], a_re_im[2 * k + 1]);
  f_ZGESVD(F_tmp<const char*>("A"), F_tmp<const char*>("A"), m, n, A, lda, S, U, lda, VT, lda, worktmp, lwork, rwork, info);
  lwork = (int)floor(worktmp[1].real());
  if (lwork > 4095) abort();
  f_ZGESVD(F_tmp<const char*>("A"), F_tmp<const char*>("A"), m, n, A, lda, S, U, lda, VT, lda, work, lwork, rwork, info);
  for (int k = 0; k < 3; ++k) s[k] = S[k + 1];
  for (int k = 0; k < 9; ++k) { u_re_im[2 * k] = U[k + 1].real(); u_re_im[2 * k + 1] = U[k + 1].imag(); vt_re_im[2 * k] = VT[k + 1].real(); vt_re_im[2 * k + 1] = VT[k + 1].imag(); }
  return info;
}
// the hook form the translated is_right_handed takes (oracle/f90_to_cpp.py: csvd_cb_t)
void lapack_ref_csvd_hook(const double* m_re_im, double* vt_re_im) {
  double s[3], u[18];
  if (lapack_ref_csvd3(m_re_im, s, u, vt_re_im) != 0) { fprintf(stderr, "lapack_ref: SVD failed\n"); abort(); }
}
}
"""

# the three calls of fortran/blas.f95 that lsinterp (fortran/lsinterp_mod.f95:375-395) makes: SolveBlasSymmetric -> dposv 'U' with one
# right-hand side, MatTransMatMultBlas -> dgemm 'T','N', MatVecMultBlas -> dgemv 'N'.  Column-major arrays, 0-based pointers.
ENTRY_DPOSV = r"""
extern "C" {
int lapack_ref_dposv_upper(int n, double* a, double* b) {
  int nrhs = 1, info = 0;
  f_DPOSV(F_tmp<const char*>("U"), n, nrhs, a - 1, n, b - 1, n, info);
  return info;
}
void lapack_ref_dgemm_tn(int m, int n, int k, double* a, double* b, double* c) {
  double one = 1.0, zero = 0.0;
  f_DGEMM(F_tmp<const char*>("T"), F_tmp<const char*>("N"), m, n, k, one, a - 1, k, b - 1, k, zero, c - 1, m);
}
void lapack_ref_dgemv_n(int m, int n, double* a, double* x, double* y) {
  double one = 1.0, zero = 0.0;
  int inc = 1;
  f_DGEMV(F_tmp<const char*>("N"), m, n, one, a - 1, m, x - 1, inc, zero, y - 1, inc);
}
}
"""


def main(dst, srcs, absent):
    ABSENT.update(absent)
    stmts = []
    for one in srcs:
        stmts += read_statements(one)
    units = parse(stmts)
    protos, texts = [], []
    for u in units.values():
        p, t = emit(u, units)
        protos.append(p)
        texts.append(t)
    stubs = "\n".join('template <class... A> int f_%s(A&&...) { fprintf(stderr, "lapack_ref: %s is on a branch a 3 x 3 call never takes\\n"); abort(); }' % (n, n)
                      for n in sorted(ABSENT))
    with open(dst, "w") as f:
        f.write(PRELUDE + "\n" + stubs + "\n\n" + "\n".join(protos) + "\n\n" + "\n".join(texts) + (ENTRY_DPOSV if "--entry=dposv" in sys.argv else ENTRY))


if __name__ == "__main__":
    absent = [a[len("--absent="):].split(",") for a in sys.argv if a.startswith("--absent=")]
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    main(args[0], args[1:], absent[0] if absent else [])
