#!/usr/bin/env python
"""A small interpreter for the array-Fortran 90 subset fortran/lsinterp_mod.f95 is written in.

TEST INFRASTRUCTURE (oracle/).  No Fortran compiler exists here, and the translators of this directory (f90_to_cpp.py) do not
cover the constructs of the scattered-sample model's interpolator (allocatable arrays of run-time shape, masks, PACK, COUNT,
vector subscripts, elemental functions applied to arrays, array constructors with implied DO).  So the reference's own source
text is EXECUTED, statement by statement, where it lies (/root/reference, read at test time only -- nothing of it is copied
into the repository): tests/test_scattered_cpu.py runs `lsinterp` this way and compares the oracle's restatement
(oracle/orc_scattered.h) with it bit for bit.

What the interpreter fixes where a compiler has a choice (the same choices as oracle/ref_stubs/f90rt.h):
  * REAL(kind=DP) arithmetic is IEEE double, one rounding per operation, evaluated left to right with the standard's
    precedence (`**` right-associative and above unary minus); array operations are element by element (numpy float64);
  * SUM and DOT_PRODUCT add in array element order (no pairwise or vector reassociation: no -ffast-math in fortran/Makefile);
  * x**2 is x*x; x**i with an integer variable is libgfortran's pow_r8_i4 (square and multiply); x**y with a real exponent,
    EXP, COS, SQRT are the C library's (glibc here, as for a gfortran binary on this machine): compare with the oracle built
    -DORC_LIBM;
  * a literal is the nearest double to its decimal text.
Derived types are objects with named components, pointers are references (NULL: None), `=>` binds a name to its target.
Procedures that are not in the interpreted files (LAPACK / BLAS; kdtree_search when only lsinterp_mod.f95 is loaded) are hooks the
caller supplies.
"""
import math
import re

import numpy as np

TOKEN = re.compile(r"\s*(?:('[^']*'|\"[^\"]*\")|(\d+\.?\d*(?:[ed][+-]?\d+)?(?:_[a-z0-9]+)?|\.\d+(?:[ed][+-]?\d+)?(?:_[a-z0-9]+)?)|(\.(?:and|or|not|true|false|eq|ne|lt|le|gt|ge)\.)"
                   r"|(\(/|/\)|\*\*|=>|==|/=|<=|>=|[-+*/(),:<>=%])|([a-z][a-z0-9_]*))")
DOTOPS = {".eq.": "==", ".ne.": "/=", ".lt.": "<", ".le.": "<=", ".gt.": ">", ".ge.": ">="}


_TOKENS = {}


def tokenize(s):
    if s not in _TOKENS:
        _TOKENS[s] = _tokenize(s)
    return list(_TOKENS[s])


def _tokenize(s):
    out, i = [], 0
    s = s.strip()
    while i < len(s):
        m = TOKEN.match(s, i)
        if not m or m.end() == i:
            raise ValueError("cannot tokenize %r at %r" % (s, s[i:]))
        string, num, dot, op, name = m.groups()
        if string is not None:
            out.append(("str", string[1:-1]))
        elif num is not None:
            # a digit string followed by a dot-operator (1.and.) does not occur in the interpreted files
            txt = re.sub(r"_[a-z0-9]+$", "", num)
            isreal = ("." in txt) or ("e" in txt) or ("d" in txt)
            out.append(("num", float(txt.replace("d", "e")) if isreal else int(txt)))
        elif dot is not None:
            out.append(("op", DOTOPS.get(dot, dot)))
        elif op is not None:
            out.append(("op", op))
        else:
            out.append(("name", name))
        i = m.end()
    return out


def read_statements(path):
    """free-form source -> list of lower-case statements, comments and continuations removed, character literals emptied"""
    stmts, cur = [], ""
    for raw in open(path):
        line, quote, buf = raw.rstrip("\n"), None, []
        for ch in line:
            if quote:
                buf.append("#" if ch in "!;&" else ch)  # (only PRINT texts hold such characters: they are never looked at)
                if ch == quote:
                    quote = None
                continue
            if ch in "'\"":
                quote = ch
                buf.append(ch)
                continue
            if ch == "!":
                break
            buf.append(ch)
        text = "".join(buf).strip().lower()
        if not text:
            continue
        if text.startswith("&"):
            text = text[1:].strip()
        if text.endswith("&"):
            cur += text[:-1] + " "
            continue
        cur += text
        for part in cur.split(";"):
            if part.strip():
                stmts.append(part.strip())
        cur = ""
    return stmts


class Unit:
    def __init__(self, kind, name, args, result):
        self.kind, self.name, self.args, self.result = kind, name, args, result
        self.body, self.types, self.dims = [], {}, {}


HEAD = re.compile(r"^(?:(?:recursive|elemental|pure)\s+)*(subroutine|function)\s+([a-z][a-z0-9_]*)\s*\(([^)]*)\)\s*(?:result\s*\(\s*([a-z][a-z0-9_]*)\s*\))?$")
DECL = re.compile(r"^(integer|real|logical|character|type)\b([^:]*)::(.*)$")


def parse_units(path):
    units, cur = {}, None
    for s in read_statements(path):
        m = HEAD.match(s)
        if m and cur is None:
            cur = Unit(m.group(1), m.group(2), [a.strip() for a in m.group(3).split(",") if a.strip()], m.group(4) or m.group(2))
            continue
        if cur is None:
            continue
        if re.match(r"^end\s*(subroutine|function)\b", s):
            units[cur.name] = cur
            cur = None
            continue
        d = DECL.match(s)
        if d and "=" not in d.group(3).split("(")[0]:
            for item in split_top(d.group(3)):
                mm = re.match(r"\s*([a-z][a-z0-9_]*)\s*(?:\((.*)\))?\s*$", item)
                cur.types[mm.group(1)] = d.group(1)
                if mm.group(2) and ":" not in mm.group(2):
                    cur.dims[mm.group(1)] = split_top(mm.group(2))  # explicit shape: automatic array, or the extent of a dummy
            continue
        if s == "implicit none":
            continue
        cur.body.append(s)
    return units


def split_top(s, sep=","):
    out, depth, cur = [], 0, ""
    i = 0
    while i < len(s):
        ch = s[i]
        if ch == "(":
            depth += 1
        elif ch == ")":
            depth -= 1
        if ch == sep and depth == 0:
            out.append(cur)
            cur = ""
        else:
            cur += ch
        i += 1
    if cur.strip():
        out.append(cur)
    return out


# ------------------------------------------------------------------ values
def _is_int(v):
    return isinstance(v, (int, np.integer)) and not isinstance(v, (bool, np.bool_)) or (isinstance(v, np.ndarray) and v.dtype.kind == "i")


def _map(f, v):
    if isinstance(v, np.ndarray):
        return np.array([f(float(t)) for t in v.ravel(order="F")], dtype=np.float64).reshape(v.shape, order="F")
    return f(float(v))


def _pow_int(x, n):  # libgfortran pow_r8_i4: square and multiply (n >= 0 here)
    n = int(n)
    if n < 0:
        return 1.0 / _pow_int(x, -n)
    pw = 1.0 if not _is_int(x) else 1
    while True:
        if n & 1:
            pw = pw * x
        n >>= 1
        if not n:
            return pw
        x = x * x


def _power(x, y):
    if _is_int(y):
        if isinstance(y, np.ndarray):
            raise NotImplementedError("array of integer exponents")
        return _pow_int(x, y)
    if isinstance(x, np.ndarray):
        return _map(lambda t: math.pow(t, float(y)), x)
    return math.pow(float(x), float(y))


def _seq_sum(a):
    a = np.asarray(a).ravel(order="F")
    if a.dtype.kind != "f":
        return int(a.sum())
    s = 0.0
    for t in a:  # array element order, one rounding per addition
        s = s + float(t)
    return s


def _divide(a, b):
    if _is_int(a) and _is_int(b):
        q = np.trunc(np.true_divide(a, b)).astype(np.int64)  # integer division truncates
        return int(q) if q.ndim == 0 else q
    return a / b if isinstance(a, np.ndarray) or isinstance(b, np.ndarray) else float(a) / float(b)


class Obj:
    """an object of a derived type: components by name (None: unallocated / disassociated)"""

    def __init__(self):
        self.attrs = {}


class Stop(Exception):
    pass


class _Return(Exception):
    pass


DTYPES = {"integer": np.int64, "real": np.float64, "logical": np.bool_}


class Interp:
    """units: name -> Unit (several files merged); consts: module parameters (pi, r_e, ...); hooks: name -> callable(values) that
    returns, for a subroutine, the list of new values of its arguments (None: unchanged) or, for a function, its value."""

    def __init__(self, units, consts, hooks):
        self.units, self.consts, self.hooks = units, dict(consts), hooks
        self.trees = {name: self._structure(u.body) for name, u in units.items()}

    # ---- statements -> nested blocks
    def _structure(self, body):
        pos = [0]

        def block(enders):
            out = []
            while pos[0] < len(body):
                s = body[pos[0]]
                key = re.sub(r"\s+", "", s)
                if any(key == e or (e.endswith("(") and key.startswith(e)) for e in enders):
                    return out
                pos[0] += 1
                m = re.match(r"^if\s*\((.*)\)\s*then$", s)
                if m:
                    arms, other = [(m.group(1), block(["elseif(", "else", "endif"]))], None
                    while True:
                        t = body[pos[0]]
                        k = re.sub(r"\s+", "", t)
                        pos[0] += 1
                        if k == "endif":
                            break
                        if k == "else":
                            other = block(["endif"])
                            continue
                        m2 = re.match(r"^else\s*if\s*\((.*)\)\s*then$", t)
                        arms.append((m2.group(1), block(["elseif(", "else", "endif"])))
                    out.append(("if", arms, other))
                    continue
                m = re.match(r"^do\s+while\s*\((.*)\)$", s)
                if m:
                    inner = block(["enddo"])
                    pos[0] += 1
                    out.append(("while", m.group(1), inner))
                    continue
                m = re.match(r"^do\s+([a-z][a-z0-9_]*)\s*=\s*(.*)$", s)
                if m:
                    inner = block(["enddo"])
                    pos[0] += 1
                    out.append(("do", m.group(1), split_top(m.group(2)), inner))
                    continue
                out.append(("stmt", s))
            return out

        return block([])

    # ---- calling
    def call(self, name, actuals):
        """actuals: list of values (None for an unallocated allocatable).  Returns (result or None, list of final argument values)."""
        u = self.units[name]
        env = {}
        for d, v in zip(u.args, actuals):
            if v is not None:
                env[d] = v
        for d in u.args:  # explicit-shape dummies see the leading elements of the actual (x(size(A,1)) in blas.f95)
            if d in u.dims and isinstance(env.get(d), np.ndarray) and env[d].ndim == 1:
                n = int(self.eval(u.dims[d][0], env, u))
                if n != env[d].shape[0]:
                    env[d] = env[d][:n]
        for nm, dims in u.dims.items():  # automatic arrays, incl. an explicit-shape function result
            if nm not in u.args:
                shape = tuple(int(self.eval(t, env, u)) for t in dims)
                kind = DTYPES.get(u.types.get(nm, "real"), np.float64)
                env[nm] = np.full(shape, np.nan, order="F") if kind is np.float64 else np.zeros(shape, dtype=kind, order="F")
        try:
            self.run(self.trees[name], env, u)
        except _Return:
            pass
        finals = [env.get(d) for d in u.args]
        return (env.get(u.result) if u.kind == "function" else None), finals

    def run(self, nodes, env, u):
        for node in nodes:
            kind = node[0]
            if kind == "if":
                done = False
                for cond, blk in node[1]:
                    if bool(self.eval(cond, env, u)):
                        self.run(blk, env, u)
                        done = True
                        break
                if not done and node[2] is not None:
                    self.run(node[2], env, u)
            elif kind == "do":
                lim = [int(self.eval(t, env, u)) for t in node[2]]
                step = lim[2] if len(lim) > 2 else 1
                i = lim[0]
                while (i <= lim[1]) if step > 0 else (i >= lim[1]):
                    env[node[1]] = i
                    self.run(node[3], env, u)
                    i += step
                env[node[1]] = i
            elif kind == "while":
                while bool(self.eval(node[1], env, u)):
                    self.run(node[2], env, u)
            else:
                self.statement(node[1], env, u)

    # ---- one simple statement
    def statement(self, s, env, u):
        key = re.sub(r"\s+", "", s)
        if key.startswith(("print*", "flush(", "write(")):
            return
        if key == "return":
            raise _Return()
        if key == "stop" or key.startswith("stop'"):
            raise Stop(u.name)
        m = re.match(r"^if\s*\(", s)
        if m:  # logical IF
            depth, j = 0, m.end() - 1
            while True:
                depth += s[j] == "("
                depth -= s[j] == ")"
                if depth == 0:
                    break
                j += 1
            if bool(self.eval(s[m.end():j], env, u)):
                self.statement(s[j + 1:].strip(), env, u)
            return
        m = re.match(r"^allocate\s*\((.*)\)$", s)
        if m:
            for item in split_top(m.group(1)):
                mm = re.match(r"\s*([a-z][a-z0-9_]*)\s*(?:%\s*([a-z][a-z0-9_]*))?\s*(?:\((.*)\))?\s*$", item)
                name, comp, dims = mm.groups()
                if dims is None:  # a pointer to an object of a derived type
                    new = Obj()
                else:
                    shape = tuple(int(self.eval(t, env, u)) for t in split_top(dims))
                    kind = DTYPES.get(u.types.get(name, "real"), np.float64) if comp is None else np.float64
                    # (allocated storage is undefined: NaN for reals, so that a use before a definition shows)
                    new = np.full(shape, np.nan, order="F") if kind is np.float64 else np.zeros(shape, dtype=kind, order="F")
                if comp is None:
                    env[name] = new
                else:
                    env[name].attrs[comp] = new
            return
        m = re.match(r"^nullify\s*\(\s*([a-z][a-z0-9_]*)\s*(?:%\s*([a-z][a-z0-9_]*))?\s*\)$", s)
        if m:
            if m.group(2):
                env[m.group(1)].attrs[m.group(2)] = None
            else:
                env.pop(m.group(1), None)
            return
        m = re.match(r"^call\s+move_alloc\s*\(\s*([a-z][a-z0-9_]*)\s*,\s*([a-z][a-z0-9_]*)\s*\)$", s)
        if m:
            env[m.group(2)] = env.pop(m.group(1))
            return
        m = re.match(r"^deallocate\s*\((.*)\)$", s)
        if m:
            for item in split_top(m.group(1)):
                env.pop(item.strip(), None)
            return
        m = re.match(r"^call\s+([a-z][a-z0-9_]*)\s*\((.*)\)$", s)
        if m:
            self.invoke(m.group(1), split_top(m.group(2)), env, u, want_value=False)
            return
        toks = tokenize(s)
        depth, eq = 0, -1
        for i, t in enumerate(toks):
            if t == ("op", "(") or t == ("op", "(/"):
                depth += 1
            elif t == ("op", ")") or t == ("op", "/)"):
                depth -= 1
            elif t == ("op", "=>") and depth == 0:  # pointer assignment: the name refers to the target itself
                env[toks[0][1]] = _Parser(self, toks[i + 1:], env, u).expression()
                return
            elif t == ("op", "=") and depth == 0:
                eq = i
                break
        if eq < 0:
            raise NotImplementedError("statement %r in %s" % (s, u.name))
        value = _Parser(self, toks[eq + 1:], env, u).expression()
        self.assign(toks[:eq], value, env, u)

    def assign(self, lhs, value, env, u):
        name = lhs[0][1]
        kind = u.types.get(name)
        if len(lhs) >= 3 and lhs[1] == ("op", "%"):
            obj, comp = env[name], lhs[2][1]
            if len(lhs) > 3:
                obj.attrs[comp][_Parser(self, lhs[4:-1], env, u).subscripts()] = value
            elif isinstance(obj.attrs.get(comp), np.ndarray) and (not isinstance(value, np.ndarray) or value.shape == obj.attrs[comp].shape):
                obj.attrs[comp][...] = value
            else:
                obj.attrs[comp] = np.array(value, order="F") if isinstance(value, np.ndarray) else value
            return
        if len(lhs) == 1:
            cur = env.get(name)
            if isinstance(cur, np.ndarray) and (not isinstance(value, np.ndarray) or value.shape == cur.shape):
                cur[...] = value
            elif isinstance(value, np.ndarray):
                env[name] = np.array(value, dtype=DTYPES.get(kind, value.dtype), order="F")  # (re)allocation on assignment
            elif kind == "integer":
                if not _is_int(value):
                    raise NotImplementedError("real assigned to integer %s" % name)
                env[name] = int(value)
            elif kind == "real":
                env[name] = float(value)
            else:
                env[name] = value
            return
        idx = _Parser(self, lhs[2:-1], env, u).subscripts()
        env[name][idx] = value

    def invoke(self, name, argtexts, env, u, want_value):
        plain = [t.strip() if re.fullmatch(r"\s*[a-z][a-z0-9_]*\s*", t) else None for t in argtexts]
        values = []
        for t, p in zip(argtexts, plain):
            if p is not None and p not in env and p not in self.consts:
                values.append(None)  # an unallocated allocatable / undefined variable handed down
            else:
                values.append(self.eval(t, env, u))
        if name in self.units:
            result, finals = self.call(name, values)
        elif name in self.hooks:
            out = self.hooks[name](values)
            result, finals = (out, [None] * len(values)) if want_value else (None, out)
        else:
            raise NotImplementedError("procedure %s" % name)
        for t, p, old, new in zip(argtexts, plain, values, finals):
            mc = re.fullmatch(r"\s*([a-z][a-z0-9_]*)\s*%\s*([a-z][a-z0-9_]*)\s*", t)
            if mc and new is not None and new is not old:
                env[mc.group(1)].attrs[mc.group(2)] = new  # e.g. a child pointer the callee allocated
                continue
            if p is None or new is None or (p in self.consts and p not in env):
                continue
            if isinstance(old, np.ndarray) and isinstance(new, np.ndarray) and new.base is old:
                continue  # the callee worked on a view of the caller's array
            if isinstance(env.get(p), np.ndarray) and isinstance(new, np.ndarray) and env[p].shape == new.shape:
                env[p][...] = new
            else:
                env[p] = new
        return result

    def eval(self, text, env, u):
        return _Parser(self, tokenize(text), env, u).expression()


class _Parser:
    def __init__(self, interp, toks, env, unit):
        self.I, self.t, self.p, self.env, self.u = interp, toks, 0, env, unit

    def peek(self):
        return self.t[self.p] if self.p < len(self.t) else (None, None)

    def take(self, op=None):
        tok = self.peek()
        if op is not None and tok != ("op", op):
            raise ValueError("expected %r, found %r in %r" % (op, tok, self.t))
        self.p += 1
        return tok

    def expression(self):
        v = self.or_()
        return v

    def or_(self):
        v = self.and_()
        while self.peek() == ("op", ".or."):
            self.take()
            w = self.and_()
            v = np.logical_or(v, w) if isinstance(v, np.ndarray) or isinstance(w, np.ndarray) else (bool(v) or bool(w))
        return v

    def and_(self):
        v = self.not_()
        while self.peek() == ("op", ".and."):
            self.take()
            w = self.not_()
            v = np.logical_and(v, w) if isinstance(v, np.ndarray) or isinstance(w, np.ndarray) else (bool(v) and bool(w))
        return v

    def not_(self):
        if self.peek() == ("op", ".not."):
            self.take()
            v = self.not_()
            return np.logical_not(v) if isinstance(v, np.ndarray) else (not bool(v))
        return self.cmp()

    def cmp(self):
        v = self.add()
        k, op = self.peek()
        if k == "op" and op in ("==", "/=", "<", "<=", ">", ">="):
            self.take()
            w = self.add()
            return {"==": lambda a, b: a == b, "/=": lambda a, b: a != b, "<": lambda a, b: a < b, "<=": lambda a, b: a <= b,
                    ">": lambda a, b: a > b, ">=": lambda a, b: a >= b}[op](v, w)
        return v

    def add(self):
        k, op = self.peek()
        if k == "op" and op in ("+", "-"):
            self.take()
            v = self.mul()
            if op == "-":
                v = -v
        else:
            v = self.mul()
        while True:
            k, op = self.peek()
            if k == "op" and op in ("+", "-"):
                self.take()
                w = self.mul()
                v = v + w if op == "+" else v - w
            else:
                return v

    def mul(self):
        v = self.pow_()
        while True:
            k, op = self.peek()
            if k == "op" and op in ("*", "/"):
                self.take()
                w = self.pow_()
                v = v * w if op == "*" else _divide(v, w)
            else:
                return v

    def pow_(self):
        v = self.primary()
        if self.peek() == ("op", "**"):
            self.take()
            w = self.pow_()  # right-associative
            v = _power(v, w)
        return v

    def subscripts(self):
        """the tokens of a subscript list (without the outer parentheses) -> numpy index, 1-based -> 0-based"""
        idx = []
        while self.p < len(self.t):
            lo = None
            if self.peek() != ("op", ":"):
                lo = self.expression()
            if self.peek() == ("op", ":"):  # a section lo:hi, either bound optional
                self.take()
                hi = None
                if self.peek()[0] is not None and self.peek() != ("op", ","):
                    hi = self.expression()
                idx.append(slice(None if lo is None else int(lo) - 1, None if hi is None else int(hi)))
            else:
                idx.append(np.asarray(lo) - 1 if isinstance(lo, np.ndarray) else int(lo) - 1)
            if self.peek() == ("op", ","):
                self.take()
            else:
                break
        return tuple(idx) if len(idx) > 1 else idx[0]

    def arg_texts(self):
        """after '(' : the token slices of the arguments up to the matching ')' (consumed)"""
        args, depth, start = [], 0, self.p
        while True:
            k, op = self.peek()
            if k is None:
                raise ValueError("unbalanced parentheses in %r" % (self.t,))
            if k == "op" and op in ("(", "(/"):
                depth += 1
            elif k == "op" and op in (")", "/)"):
                if depth == 0:
                    if self.p > start:
                        args.append(self.t[start:self.p])
                    self.take()
                    return args
                depth -= 1
            elif k == "op" and op == "," and depth == 0:
                args.append(self.t[start:self.p])
                start = self.p + 1
            self.p += 1

    def sub(self, toks):
        return _Parser(self.I, toks, self.env, self.u)

    def primary(self):
        k, v = self.take()
        if k == "num" or k == "str":
            return v
        if k == "op" and v == ".true.":
            return True
        if k == "op" and v == ".false.":
            return False
        if k == "op" and v == "(":
            r = self.expression()
            self.take(")")
            return r
        if k == "op" and v == "(/":
            return self.constructor()
        if k != "name":
            raise ValueError("unexpected %r in %r" % ((k, v), self.t))
        if self.peek() == ("op", "%"):  # components of an object of a derived type
            obj = self.env.get(v)
            while self.peek() == ("op", "%"):
                self.take()
                comp = self.take()[1]
                obj = obj.attrs.get(comp) if obj is not None else None
            if self.peek() == ("op", "(") and isinstance(obj, np.ndarray):
                self.take("(")
                return self.index(obj)
            return obj
        if self.peek() != ("op", "("):
            if v in self.env:
                return self.env[v]
            if v in self.I.consts:
                return self.I.consts[v]
            raise NameError("%s is used before it is defined (in %s)" % (v, self.u.name))
        self.take("(")
        if isinstance(self.env.get(v), np.ndarray):
            return self.index(self.env[v])
        args = self.arg_texts()
        return self.function(v, args)

    def index(self, arr):
        """after '(' of a subscript list on `arr`"""
        args = self.arg_texts()
        flat = []
        for i, a in enumerate(args):
            flat += list(a) + ([("op", ",")] if i + 1 < len(args) else [])
        got = arr[self.sub(flat).subscripts()]
        return got.item() if isinstance(got, np.generic) or (isinstance(got, np.ndarray) and got.ndim == 0) else got

    def designator(self):
        """value of a variable that may be unallocated / disassociated / undefined: None then"""
        try:
            return self.expression()
        except NameError:
            return None

    def constructor(self):
        items = []
        while self.peek() != ("op", "/)"):
            if self.peek() == ("op", "("):  # implied DO: ( expr, var = lo, hi )
                save = self.p
                self.take("(")
                parts = self.arg_texts()
                eqs = [i for i, a in enumerate(parts) if len(a) > 2 and a[0][0] == "name" and a[1] == ("op", "=")]
                if eqs:
                    e = eqs[0]
                    var, lo, hi = parts[e][0][1], int(self.sub(parts[e][2:]).expression()), int(self.sub(parts[e + 1]).expression())
                    for i in range(lo, hi + 1):
                        self.env[var] = i
                        for a in parts[:e]:
                            items.append(self.sub(a).expression())
                    self.env[var] = hi + 1
                else:
                    self.p = save
                    items.append(self.expression())
            else:
                items.append(self.expression())
            if self.peek() == ("op", ","):
                self.take()
        self.take("/)")
        if all(_is_int(i) for i in items):
            return np.array(items, dtype=np.int64)
        return np.array(items, dtype=np.float64)

    def function(self, name, args):
        ev = lambda a: self.sub(a).expression()
        if name == "allocated":
            return isinstance(self.sub(args[0]).designator(), np.ndarray)
        if name == "associated":
            return self.sub(args[0]).designator() is not None
        if name == "size":
            a = ev(args[0])
            return int(a.size) if len(args) == 1 else int(a.shape[int(ev(args[1])) - 1])
        if name == "sum":
            return _seq_sum(ev(args[0]))
        if name == "dot_product":
            return _seq_sum(ev(args[0]) * ev(args[1]))
        if name == "count":
            return int(np.count_nonzero(ev(args[0])))
        if name == "pack":
            return np.array(ev(args[0])[np.asarray(ev(args[1]), dtype=bool)], order="F")
        if name in ("sqrt", "exp", "cos", "sin", "log"):
            return _map({"sqrt": math.sqrt, "exp": math.exp, "cos": math.cos, "sin": math.sin, "log": math.log}[name], ev(args[0]))
        if name == "floor":
            return int(math.floor(ev(args[0])))
        if name == "real":
            return float(ev(args[0]))
        if name == "mod":
            return int(math.fmod(ev(args[0]), ev(args[1])))
        texts = [" ".join(_untok(t) for t in a) for a in args]
        return self.I.invoke(name, texts, self.env, self.u, want_value=True)


def _untok(t):
    k, v = t
    if k == "str":
        return "'%s'" % v
    if k == "num":
        return repr(v)
    return v
