"""jsmini: a minimal JavaScript-subset interpreter that EXECUTES the reference's own source lines.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Why: the reference (a WeChat mini-program) ships no tests and no golden vectors, and this image has no
JavaScript engine (node, deno, quickjs ... all absent).  The layout arithmetic of the hot
path is nevertheless a tiny, self-contained subset of JavaScript -- let/const, arithmetic on IEEE doubles,
ternaries, if/while/for/switch, Math.*, Number.isFinite, Array.prototype.map/filter/reduce with arrow
functions, `||`/`&&`, template literals -- so this module tokenises, parses (Pratt) and tree-walks exactly
that subset and runs the VERBATIM text of

    pages/index/index.js:1193, 1581, 1611-1615   the stitchProgress values 1, 96, 100
    pages/index/index.js:1211-1213   totalBytes, bigTask, noCacheInThisStitch
    pages/index/index.js:1243-1244, 1246-1248   dimension normalisation, probe progress
    pages/index/index.js:1251-1386   base size, integerise, limits, clamp, supersample, guard loop
    pages/index/index.js:1423-1433   white fill, ctx.scale(superSample), scaledGap, cursors
    pages/index/index.js:1522-1557   per-image dest rect, drawWithOrientation call, cursor, progress
    utils/canvas.js:153-202          drawWithOrientation (CTM per EXIF orientation)

sliced by line number out of /root/reference, under a ~20-line `this` / `wx` / `console` / ctx stub
(`run_reference_layout`).  Numbers are Python floats (= JS numbers); Math.round is ECMA-262's (ties to +inf).
The interpreter knows nothing about stitching: every value it returns was computed by the reference's text.
`oracle/gen_golden.py` regenerates tests/golden/layout_golden.json from these runs, and tests/test_jsmini.py
holds the restatement (layout_ref.py), the C++ twin and the CTM table against them.

/root/reference exists only in the build container: nothing under `-m gpu`, smoke() or bench.py imports this.
"""
from __future__ import annotations

import math
import os
import re

REF_ROOT = "/root/reference/miniprogram-stitch/miniprogram"
INDEX_JS = os.path.join(REF_ROOT, "pages", "index", "index.js")
CANVAS_JS = os.path.join(REF_ROOT, "utils", "canvas.js")


# ----------------------------------------------------------------------------------------------------------
# values
# ----------------------------------------------------------------------------------------------------------
class _Undefined:
    def __repr__(self):
        return "undefined"

    def __bool__(self):
        return False


UNDEF = _Undefined()


class JSError(Exception):
    """A JavaScript `throw`: .value is the thrown value (an Error object is a dict with 'message')."""

    def __init__(self, value):
        super().__init__(value.get("message") if isinstance(value, dict) else str(value))
        self.value = value


class _Break(Exception):
    pass


class _Continue(Exception):
    pass


class _Return(Exception):
    def __init__(self, value):
        self.value = value


def truthy(v):
    if v is UNDEF or v is None or v is False:
        return False
    if v is True:
        return True
    if isinstance(v, float):
        return not (v == 0.0 or v != v)
    if isinstance(v, str):
        return len(v) > 0
    return True


def js_round(x):
    if x != x or x in (math.inf, -math.inf):
        return x
    f = math.floor(x)
    return float(f + 1) if (x - f) >= 0.5 else float(f)


def to_number(v):
    if isinstance(v, bool):
        return 1.0 if v else 0.0
    if isinstance(v, float):
        return v
    if v is None:
        return 0.0
    if v is UNDEF:
        return math.nan
    if isinstance(v, str):
        try:
            return float(v) if v.strip() else 0.0
        except ValueError:
            return math.nan
    return math.nan


def num_to_string(x):
    if isinstance(x, float):
        if x != x:
            return "NaN"
        if x in (math.inf, -math.inf):
            return "Infinity" if x > 0 else "-Infinity"
        if x == math.floor(x) and abs(x) < 1e21:
            return str(int(x))
        return repr(x)
    return to_string(x)


def to_string(v):
    if isinstance(v, str):
        return v
    if isinstance(v, bool):
        return "true" if v else "false"
    if isinstance(v, float):
        return num_to_string(v)
    if v is None:
        return "null"
    if v is UNDEF:
        return "undefined"
    if isinstance(v, list):
        return ",".join(to_string(x) for x in v)
    return "[object Object]"


class NativeFn:
    """A host function; `apply(thisArg, argsArray)` and `call` are provided as Function.prototype does."""

    def __init__(self, fn, name="native"):
        self.fn = fn
        self.name = name

    def __call__(self, this, args):
        return self.fn(this, args)


class Closure:
    def __init__(self, params, body, env, this, is_arrow, is_expr):
        self.params, self.body, self.env, self.this = params, body, env, this
        self.is_arrow, self.is_expr = is_arrow, is_expr


# ----------------------------------------------------------------------------------------------------------
# tokenizer
# ----------------------------------------------------------------------------------------------------------
_PUNCT = ["===", "!==", "=>", "...", "==", "!=", "<=", ">=", "&&", "||", "++", "--", "+=", "-=", "*=", "/=",
          "{", "}", "(", ")", "[", "]", ";", ",", "<", ">", "+", "-", "*", "/", "%", "!", "?", ":", "=", "."]
_KEYWORDS = {"let", "const", "var", "if", "else", "for", "while", "break", "continue", "return", "throw", "new", "function",
             "switch", "case", "default", "try", "catch", "finally", "true", "false", "null", "undefined", "typeof", "this",
             "await", "async"}
_NUM = re.compile(r"0[xX][0-9a-fA-F]+|(?:\d+\.?\d*|\.\d+)(?:[eE][+-]?\d+)?")
_ID = re.compile(r"[A-Za-z_$][A-Za-z0-9_$]*")


def tokenize(src):
    toks = []   # (kind, value); kinds: num, str, tpl, id, kw, p, eof
    i, n = 0, len(src)
    while i < n:
        c = src[i]
        if c in " \t\r\n":
            i += 1
            continue
        if src.startswith("//", i):
            j = src.find("\n", i)
            i = n if j < 0 else j
            continue
        if src.startswith("/*", i):
            j = src.find("*/", i + 2)
            i = n if j < 0 else j + 2
            continue
        if c in "'\"":
            j, out = i + 1, []
            while src[j] != c:
                if src[j] == "\\":
                    j += 1
                    out.append({"n": "\n", "t": "\t"}.get(src[j], src[j]))
                else:
                    out.append(src[j])
                j += 1
            toks.append(("str", "".join(out)))
            i = j + 1
            continue
        if c == "`":   # template literal -> ('tpl', [str | ('expr', source)])
            j, parts, cur = i + 1, [], []
            while src[j] != "`":
                if src.startswith("${", j):
                    depth, k = 1, j + 2
                    while depth:
                        depth += {"{": 1, "}": -1}.get(src[k], 0)
                        k += 1
                    parts.append("".join(cur))
                    cur = []
                    parts.append(("expr", src[j + 2:k - 1]))
                    j = k
                else:
                    cur.append(src[j])
                    j += 1
            parts.append("".join(cur))
            toks.append(("tpl", parts))
            i = j + 1
            continue
        m = _NUM.match(src, i)
        if m and (c.isdigit() or (c == "." and i + 1 < n and src[i + 1].isdigit())):
            t = m.group(0)
            toks.append(("num", float(int(t, 16)) if t[:2].lower() == "0x" else float(t)))
            i = m.end()
            continue
        m = _ID.match(src, i)
        if m:
            t = m.group(0)
            toks.append(("kw" if t in _KEYWORDS else "id", t))
            i = m.end()
            continue
        for p in _PUNCT:
            if src.startswith(p, i):
                toks.append(("p", p))
                i += len(p)
                break
        else:
            raise SyntaxError(f"jsmini: unexpected character {c!r} at {i}: {src[max(0, i - 30):i + 30]!r}")
    toks.append(("eof", None))
    return toks


# ----------------------------------------------------------------------------------------------------------
# parser -> AST of tuples
# ----------------------------------------------------------------------------------------------------------
_BINARY = {"||": 1, "&&": 2, "===": 5, "!==": 5, "==": 5, "!=": 5, "<": 6, ">": 6, "<=": 6, ">=": 6,
           "+": 8, "-": 8, "*": 9, "/": 9, "%": 9}
_ASSIGN = {"=", "+=", "-=", "*=", "/="}


class Parser:
    def __init__(self, src):
        self.t = tokenize(src)
        self.i = 0

    # -- token helpers
    def peek(self, k=0):
        return self.t[self.i + k]

    def at(self, kind, val=None, k=0):
        t = self.t[self.i + k]
        return t[0] == kind and (val is None or t[1] == val)

    def eat(self, kind, val=None):
        if self.at(kind, val):
            self.i += 1
            return True
        return False

    def expect(self, kind, val=None):
        if not self.at(kind, val):
            ctx = " ".join(str(x[1]) for x in self.t[max(0, self.i - 8):self.i + 4])
            raise SyntaxError(f"jsmini: expected {val or kind}, got {self.peek()} near: {ctx}")
        tok = self.t[self.i]
        self.i += 1
        return tok

    # -- statements
    def program(self):
        body = []
        while not self.at("eof"):
            body.append(self.statement())
        return ("block", body)

    def block(self):
        self.expect("p", "{")
        body = []
        while not self.at("p", "}"):
            body.append(self.statement())
        self.expect("p", "}")
        return ("block", body)

    def statement(self):
        if self.at("p", "{"):
            return self.block()
        if self.at("p", ";"):
            self.i += 1
            return ("empty",)
        if self.at("kw"):
            kw = self.peek()[1]
            if kw in ("let", "const", "var"):
                self.i += 1
                decls = []
                while True:
                    name = self.expect("id")[1]
                    init = self.assignment() if self.eat("p", "=") else None
                    decls.append((name, init))
                    if not self.eat("p", ","):
                        break
                self.eat("p", ";")
                return ("decl", decls)
            if kw == "if":
                self.i += 1
                self.expect("p", "(")
                c = self.expression()
                self.expect("p", ")")
                a = self.statement()
                b = self.statement() if self.eat("kw", "else") else None
                return ("if", c, a, b)
            if kw == "while":
                self.i += 1
                self.expect("p", "(")
                c = self.expression()
                self.expect("p", ")")
                return ("while", c, self.statement())
            if kw == "for":
                self.i += 1
                self.expect("p", "(")
                init = None if self.at("p", ";") else self.statement_no_semicolon()
                self.expect("p", ";")
                cond = None if self.at("p", ";") else self.expression()
                self.expect("p", ";")
                step = None if self.at("p", ")") else self.expression()
                self.expect("p", ")")
                return ("for", init, cond, step, self.statement())
            if kw == "break":
                self.i += 1
                self.eat("p", ";")
                return ("break",)
            if kw == "continue":
                self.i += 1
                self.eat("p", ";")
                return ("continue",)
            if kw == "return":
                self.i += 1
                v = None if (self.at("p", ";") or self.at("p", "}")) else self.expression()
                self.eat("p", ";")
                return ("return", v)
            if kw == "throw":
                self.i += 1
                v = self.expression()
                self.eat("p", ";")
                return ("throw", v)
            if kw == "switch":
                self.i += 1
                self.expect("p", "(")
                d = self.expression()
                self.expect("p", ")")
                self.expect("p", "{")
                cases = []
                while not self.at("p", "}"):
                    if self.eat("kw", "default"):
                        test = None
                    else:
                        self.expect("kw", "case")
                        test = self.expression()
                    self.expect("p", ":")
                    body = []
                    while not (self.at("kw", "case") or self.at("kw", "default") or self.at("p", "}")):
                        body.append(self.statement())
                    cases.append((test, body))
                self.expect("p", "}")
                return ("switch", d, cases)
            if kw == "try":
                self.i += 1
                body = self.block()
                param, handler, final = None, None, None
                if self.eat("kw", "catch"):
                    if self.eat("p", "("):
                        param = self.expect("id")[1]
                        self.expect("p", ")")
                    handler = self.block()
                if self.eat("kw", "finally"):
                    final = self.block()
                return ("try", body, param, handler, final)
            if kw == "function" and self.at("id", None, 1):
                self.i += 1
                name = self.expect("id")[1]
                params = self.params()
                return ("decl", [(name, ("func", params, self.block(), False, False))])
        e = self.expression()
        self.eat("p", ";")
        return ("expr", e)

    def statement_no_semicolon(self):
        if self.at("kw") and self.peek()[1] in ("let", "const", "var"):
            self.i += 1
            decls = []
            while True:
                name = self.expect("id")[1]
                init = self.assignment() if self.eat("p", "=") else None
                decls.append((name, init))
                if not self.eat("p", ","):
                    break
            return ("decl", decls)
        return ("expr", self.expression())

    def params(self):
        self.expect("p", "(")
        ps = []
        while not self.at("p", ")"):
            ps.append(self.expect("id")[1])
            self.eat("p", ",")
        self.expect("p", ")")
        return ps

    # -- expressions
    def expression(self):
        e = self.assignment()
        while self.eat("p", ","):
            e = ("seq", e, self.assignment())
        return e

    def is_arrow_ahead(self):
        if self.at("id") and self.at("p", "=>", 1):
            return True
        if not self.at("p", "("):
            return False
        depth, k = 0, self.i
        while True:
            t = self.t[k]
            if t == ("p", "("):
                depth += 1
            elif t == ("p", ")"):
                depth -= 1
                if depth == 0:
                    return self.t[k + 1] == ("p", "=>")
            elif t[0] == "eof":
                return False
            k += 1

    def assignment(self):
        if self.at("kw", "async") and (self.at("p", "(", 1) or self.at("id", None, 1)):
            self.i += 1
        if self.is_arrow_ahead():
            params = [self.expect("id")[1]] if self.at("id") else self.params()
            self.expect("p", "=>")
            if self.at("p", "{"):
                return ("func", params, self.block(), True, False)
            return ("func", params, self.assignment(), True, True)
        left = self.ternary()
        if self.at("p") and self.peek()[1] in _ASSIGN:
            op = self.peek()[1]
            self.i += 1
            return ("assign", op, left, self.assignment())
        return left

    def ternary(self):
        c = self.binary(1)
        if self.eat("p", "?"):
            a = self.assignment()
            self.expect("p", ":")
            return ("cond", c, a, self.assignment())
        return c

    def binary(self, min_prec):
        left = self.unary()
        while self.at("p") and self.peek()[1] in _BINARY and _BINARY[self.peek()[1]] >= min_prec:
            op = self.peek()[1]
            self.i += 1
            right = self.binary(_BINARY[op] + 1)
            left = ("logic" if op in ("||", "&&") else "bin", op, left, right)
        return left

    def unary(self):
        if self.at("p") and self.peek()[1] in ("!", "-", "+"):
            op = self.peek()[1]
            self.i += 1
            return ("unary", op, self.unary())
        if self.at("p") and self.peek()[1] in ("++", "--"):
            op = self.peek()[1]
            self.i += 1
            return ("update", op, True, self.unary())
        if self.at("kw", "typeof"):
            self.i += 1
            return ("typeof", self.unary())
        if self.at("kw", "await"):
            self.i += 1
            return self.unary()   # promises are not modelled: every stubbed call is synchronous
        return self.postfix()

    def postfix(self):
        e = self.call_member()
        if self.at("p") and self.peek()[1] in ("++", "--"):
            op = self.peek()[1]
            self.i += 1
            return ("update", op, False, e)
        return e

    def args(self):
        self.expect("p", "(")
        a = []
        while not self.at("p", ")"):
            a.append(self.assignment())
            self.eat("p", ",")
        self.expect("p", ")")
        return a

    def call_member(self):
        if self.at("kw", "new"):
            self.i += 1
            callee = self.primary()
            while self.eat("p", "."):
                callee = ("member", callee, ("lit", self.t[self.i][1]))
                self.i += 1
            e = ("new", callee, self.args() if self.at("p", "(") else [])
        else:
            e = self.primary()
        while True:
            if self.eat("p", "."):
                name = self.t[self.i][1]
                self.i += 1
                e = ("member", e, ("lit", name))
            elif self.at("p", "["):
                self.i += 1
                idx = self.expression()
                self.expect("p", "]")
                e = ("member", e, idx)
            elif self.at("p", "("):
                e = ("call", e, self.args())
            else:
                return e

    def primary(self):
        k, v = self.peek()
        if k == "num" or k == "str":
            self.i += 1
            return ("lit", v)
        if k == "tpl":
            self.i += 1
            return ("tpl", [p if isinstance(p, str) else Parser(p[1]).expression() for p in v])
        if k == "id":
            self.i += 1
            return ("name", v)
        if k == "kw":
            if v in ("true", "false"):
                self.i += 1
                return ("lit", v == "true")
            if v == "null":
                self.i += 1
                return ("lit", None)
            if v == "undefined":
                self.i += 1
                return ("lit", UNDEF)
            if v == "this":
                self.i += 1
                return ("this",)
            if v == "function":
                self.i += 1
                if self.at("id"):
                    self.i += 1
                params = self.params()
                return ("func", params, self.block(), False, False)
        if k == "p" and v == "(":
            self.i += 1
            e = self.expression()
            self.expect("p", ")")
            return e
        if k == "p" and v == "[":
            self.i += 1
            items = []
            while not self.at("p", "]"):
                items.append(self.assignment())
                self.eat("p", ",")
            self.expect("p", "]")
            return ("array", items)
        if k == "p" and v == "{":
            self.i += 1
            props = []
            while not self.at("p", "}"):
                if self.eat("p", "..."):
                    props.append(("spread", self.assignment()))
                else:
                    key = self.t[self.i][1]
                    self.i += 1
                    if self.eat("p", ":"):
                        props.append((key, self.assignment()))
                    else:
                        props.append((key, ("name", key)))   # shorthand
                self.eat("p", ",")
            self.expect("p", "}")
            return ("object", props)
        raise SyntaxError(f"jsmini: unexpected token {self.peek()} near: " +
                          " ".join(str(x[1]) for x in self.t[max(0, self.i - 8):self.i + 4]))


# ----------------------------------------------------------------------------------------------------------
# evaluator
# ----------------------------------------------------------------------------------------------------------
class Env:
    def __init__(self, parent=None):
        self.vars = {}
        self.parent = parent

    def lookup(self, name):
        e = self
        while e is not None:
            if name in e.vars:
                return e
            e = e.parent
        return None

    def get(self, name):
        e = self.lookup(name)
        if e is None:
            raise JSError({"message": f"ReferenceError: {name} is not defined"})
        return e.vars[name]

    def set(self, name, value):
        e = self.lookup(name)
        if e is None:
            raise JSError({"message": f"ReferenceError: {name} is not defined"})
        e.vars[name] = value

    def declare(self, name, value):
        self.vars[name] = value


def _arr_method(arr, name, interp):
    def call_fn(f, args):
        return interp.call(f, UNDEF, args)

    if name == "map":
        return NativeFn(lambda this, a: [call_fn(a[0], [x, float(i), arr]) for i, x in enumerate(arr)])
    if name == "filter":
        return NativeFn(lambda this, a: [x for i, x in enumerate(arr) if truthy(call_fn(a[0], [x, float(i), arr]))])
    if name == "forEach":
        def each(this, a):
            for i, x in enumerate(arr):
                call_fn(a[0], [x, float(i), arr])
            return UNDEF
        return NativeFn(each)
    if name == "reduce":
        def reduce(this, a):
            it = list(enumerate(arr))
            if len(a) > 1:
                acc = a[1]
            else:
                acc = it.pop(0)[1]
            for i, x in it:
                acc = call_fn(a[0], [acc, x, float(i), arr])
            return acc
        return NativeFn(reduce)
    if name == "push":
        def push(this, a):
            arr.extend(a)
            return float(len(arr))
        return NativeFn(push)
    if name == "slice":
        return NativeFn(lambda this, a: arr[int(a[0]) if a else 0:int(a[1]) if len(a) > 1 else None])
    return UNDEF


def _to_fixed(x, digits):
    d = int(digits)
    # JS toFixed rounds half up on the exact decimal expansion; Python's format rounds half even on the same
    # expansion -- they differ only on exact ties, which only affect console.log text here.
    return f"{x:.{d}f}"


class Interp:
    def __init__(self):
        self.globals = Env()
        g = self.globals
        math_obj = {
            "PI": math.pi,
            "round": NativeFn(lambda t, a: js_round(to_number(a[0]))),
            "floor": NativeFn(lambda t, a: float(math.floor(to_number(a[0]))) if math.isfinite(to_number(a[0])) else to_number(a[0])),
            "ceil": NativeFn(lambda t, a: float(math.ceil(to_number(a[0]))) if math.isfinite(to_number(a[0])) else to_number(a[0])),
            "sqrt": NativeFn(lambda t, a: math.sqrt(to_number(a[0])) if to_number(a[0]) >= 0 else math.nan),
            "abs": NativeFn(lambda t, a: abs(to_number(a[0]))),
            "min": NativeFn(lambda t, a: self._minmax(a, min, math.inf)),
            "max": NativeFn(lambda t, a: self._minmax(a, max, -math.inf)),
        }
        g.declare("Math", math_obj)
        g.declare("Number", {"isFinite": NativeFn(lambda t, a: isinstance(a[0], float) and math.isfinite(a[0]))})
        g.declare("Error", NativeFn(lambda t, a: {"message": to_string(a[0]) if a else ""}, "Error"))
        g.declare("console", {k: NativeFn(lambda t, a: UNDEF) for k in ("log", "warn", "error", "info")})

    @staticmethod
    def _minmax(args, fn, empty):
        vals = [to_number(x) for x in args]
        if not vals:
            return empty
        if any(v != v for v in vals):
            return math.nan
        return float(fn(vals))

    # -- calls
    def call(self, f, this, args):
        if isinstance(f, NativeFn):
            return f(this, args)
        if isinstance(f, Closure):
            env = Env(f.env)
            for i, p in enumerate(f.params):
                env.declare(p, args[i] if i < len(args) else UNDEF)
            this_val = f.this if f.is_arrow else this
            if f.is_expr:
                return self.eval(f.body, env, this_val)
            try:
                self.exec(f.body, env, this_val)
            except _Return as r:
                return r.value
            return UNDEF
        raise JSError({"message": f"TypeError: {f!r} is not a function"})

    def get_member(self, obj, key):
        if isinstance(obj, list):
            if key == "length":
                return float(len(obj))
            if isinstance(key, float):
                i = int(key)
                return obj[i] if 0 <= i < len(obj) else UNDEF
            return _arr_method(obj, key, self)
        if isinstance(obj, dict):
            return obj.get(key, UNDEF)
        if isinstance(obj, (NativeFn, Closure)):
            if key == "apply":
                return NativeFn(lambda t, a: self.call(obj, a[0] if a else UNDEF, list(a[1]) if len(a) > 1 and a[1] else []))
            if key == "call":
                return NativeFn(lambda t, a: self.call(obj, a[0] if a else UNDEF, list(a[1:])))
            return UNDEF
        if isinstance(obj, float):
            if key == "toFixed":
                return NativeFn(lambda t, a: _to_fixed(obj, a[0] if a else 0.0))
            if key == "toString":
                return NativeFn(lambda t, a: num_to_string(obj))
            return UNDEF
        if isinstance(obj, str):
            if key == "length":
                return float(len(obj))
            return UNDEF
        if hasattr(obj, "js_get"):
            return obj.js_get(key)
        if obj is UNDEF or obj is None:
            raise JSError({"message": f"TypeError: cannot read property {key!r} of {to_string(obj)}"})
        return UNDEF

    def set_member(self, obj, key, value):
        if isinstance(obj, dict):
            obj[key] = value
        elif isinstance(obj, list) and isinstance(key, float):
            i = int(key)
            while len(obj) <= i:
                obj.append(UNDEF)
            obj[i] = value
        elif hasattr(obj, "js_set"):
            obj.js_set(key, value)
        else:
            raise JSError({"message": f"TypeError: cannot set property {key!r}"})

    # -- operators
    @staticmethod
    def strict_equal(a, b):
        if isinstance(a, float) and isinstance(b, float):
            return a == b
        if type(a) is not type(b):
            return False
        if isinstance(a, (str, bool)):
            return a == b
        return a is b

    def binop(self, op, a, b):
        if op == "+":
            if isinstance(a, str) or isinstance(b, str) or isinstance(a, (dict, list)) or isinstance(b, (dict, list)):
                return to_string(a) + to_string(b)
            return to_number(a) + to_number(b)
        if op in ("===", "!=="):
            r = self.strict_equal(a, b)
            return r if op == "===" else not r
        if op in ("==", "!="):
            if (a is None or a is UNDEF) and (b is None or b is UNDEF):
                r = True
            elif isinstance(a, str) and isinstance(b, str):
                r = a == b
            elif a is None or a is UNDEF or b is None or b is UNDEF:
                r = False
            else:
                r = to_number(a) == to_number(b)
            return r if op == "==" else not r
        if op in ("<", ">", "<=", ">="):
            if isinstance(a, str) and isinstance(b, str):
                x, y = a, b
            else:
                x, y = to_number(a), to_number(b)
                if x != x or y != y:
                    return False
            return {"<": x < y, ">": x > y, "<=": x <= y, ">=": x >= y}[op]
        x, y = to_number(a), to_number(b)
        if op == "-":
            return x - y
        if op == "*":
            if (x in (math.inf, -math.inf) and y == 0) or (y in (math.inf, -math.inf) and x == 0):
                return math.nan
            return x * y
        if op == "/":
            if y == 0:
                if x == 0 or x != x:
                    return math.nan
                neg = (x < 0) != (math.copysign(1.0, y) < 0)
                return -math.inf if neg else math.inf
            return x / y
        if op == "%":
            if y == 0 or x in (math.inf, -math.inf) or x != x or y != y:
                return math.nan
            return math.fmod(x, y)
        raise NotImplementedError(op)

    # -- expressions
    def eval(self, n, env, this):
        k = n[0]
        if k == "lit":
            return n[1]
        if k == "name":
            return env.get(n[1])
        if k == "this":
            return this
        if k == "tpl":
            return "".join(p if isinstance(p, str) else to_string(self.eval(p, env, this)) for p in n[1])
        if k == "member":
            obj = self.eval(n[1], env, this)
            key = self.eval(n[2], env, this)
            return self.get_member(obj, key)
        if k == "call":
            callee = n[1]
            if callee[0] == "member":
                obj = self.eval(callee[1], env, this)
                f = self.get_member(obj, self.eval(callee[2], env, this))
                this_arg = obj
            else:
                f = self.eval(callee, env, this)
                this_arg = UNDEF
            args = [self.eval(a, env, this) for a in n[2]]
            if f is UNDEF:
                raise JSError({"message": f"TypeError: {callee} is not a function"})
            return self.call(f, this_arg, args)
        if k == "new":
            f = self.eval(n[1], env, this)
            return self.call(f, {}, [self.eval(a, env, this) for a in n[2]])
        if k == "bin":
            return self.binop(n[1], self.eval(n[2], env, this), self.eval(n[3], env, this))
        if k == "logic":
            a = self.eval(n[2], env, this)
            if n[1] == "||":
                return a if truthy(a) else self.eval(n[3], env, this)
            return self.eval(n[3], env, this) if truthy(a) else a
        if k == "unary":
            v = self.eval(n[2], env, this)
            if n[1] == "!":
                return not truthy(v)
            return -to_number(v) if n[1] == "-" else to_number(v)
        if k == "typeof":
            try:
                v = self.eval(n[1], env, this)
            except JSError:
                return "undefined"
            if v is UNDEF:
                return "undefined"
            if isinstance(v, bool):
                return "boolean"
            if isinstance(v, float):
                return "number"
            if isinstance(v, str):
                return "string"
            if isinstance(v, (NativeFn, Closure)):
                return "function"
            return "object"
        if k == "cond":
            return self.eval(n[2] if truthy(self.eval(n[1], env, this)) else n[3], env, this)
        if k == "assign":
            op, target, rhs = n[1], n[2], n[3]
            if op == "=":
                v = self.eval(rhs, env, this)
            else:
                v = self.binop(op[0], self.eval(target, env, this), self.eval(rhs, env, this))
            self.store(target, v, env, this)
            return v
        if k == "update":
            old = to_number(self.eval(n[3], env, this))
            new = old + (1.0 if n[1] == "++" else -1.0)
            self.store(n[3], new, env, this)
            return new if n[2] else old
        if k == "func":
            return Closure(n[1], n[2], env, this, n[3], n[4])
        if k == "array":
            return [self.eval(x, env, this) for x in n[1]]
        if k == "object":
            o = {}
            for key, val in n[1]:
                if key == "spread":
                    src = self.eval(val, env, this)
                    if isinstance(src, dict):
                        o.update(src)
                else:
                    o[key] = self.eval(val, env, this)
            return o
        if k == "seq":
            self.eval(n[1], env, this)
            return self.eval(n[2], env, this)
        raise NotImplementedError(k)

    def store(self, target, v, env, this):
        if target[0] == "name":
            env.set(target[1], v)
        elif target[0] == "member":
            self.set_member(self.eval(target[1], env, this), self.eval(target[2], env, this), v)
        else:
            raise JSError({"message": "SyntaxError: invalid assignment target"})

    # -- statements
    def exec(self, n, env, this):
        k = n[0]
        if k == "block":
            inner = Env(env)
            for s in n[1]:
                self.exec(s, inner, this)
        elif k == "decl":
            for name, init in n[1]:
                env.declare(name, self.eval(init, env, this) if init is not None else UNDEF)
        elif k == "expr":
            self.eval(n[1], env, this)
        elif k == "if":
            if truthy(self.eval(n[1], env, this)):
                self.exec(n[2], env, this)
            elif n[3] is not None:
                self.exec(n[3], env, this)
        elif k == "while":
            while truthy(self.eval(n[1], env, this)):
                try:
                    self.exec(n[2], env, this)
                except _Break:
                    break
                except _Continue:
                    continue
        elif k == "for":
            loop = Env(env)
            if n[1] is not None:
                self.exec(n[1], loop, this)
            while n[2] is None or truthy(self.eval(n[2], loop, this)):
                try:
                    self.exec(n[4], loop, this)
                except _Break:
                    break
                except _Continue:
                    pass
                if n[3] is not None:
                    self.eval(n[3], loop, this)
        elif k == "break":
            raise _Break()
        elif k == "continue":
            raise _Continue()
        elif k == "return":
            raise _Return(self.eval(n[1], env, this) if n[1] is not None else UNDEF)
        elif k == "throw":
            raise JSError(self.eval(n[1], env, this))
        elif k == "switch":
            d = self.eval(n[1], env, this)
            inner = Env(env)
            matched = False
            try:
                for test, body in n[2]:
                    if not matched and test is not None and self.strict_equal(d, self.eval(test, inner, this)):
                        matched = True
                    if matched:
                        for s in body:
                            self.exec(s, inner, this)
                if not matched:
                    run = False
                    for test, body in n[2]:
                        run = run or test is None
                        if run:
                            for s in body:
                                self.exec(s, inner, this)
            except _Break:
                pass
        elif k == "try":
            try:
                try:
                    self.exec(n[1], env, this)
                except JSError as e:
                    if n[3] is None:
                        raise
                    h = Env(env)
                    if n[2]:
                        h.declare(n[2], e.value)
                    self.exec(n[3], h, this)
            finally:
                if n[4] is not None:
                    self.exec(n[4], env, this)
        elif k == "empty":
            pass
        else:
            raise NotImplementedError(k)

    def run(self, src, env=None, this=UNDEF):
        self.exec_top(Parser(src).program(), env or self.globals, this)

    def exec_top(self, prog, env, this):
        for s in prog[1]:   # top level: declarations land in `env` itself
            self.exec(s, env, this)


# ----------------------------------------------------------------------------------------------------------
# the harness: the reference's verbatim lines under a minimal page / wx / canvas stub
# ----------------------------------------------------------------------------------------------------------
def ref_lines(path, first, last):
    """Lines first..last (1-based, inclusive) of a reference file, verbatim."""
    with open(path, encoding="utf-8") as f:
        lines = f.read().split("\n")
    return "\n".join(lines[first - 1:last])


def reference_available():
    return os.path.exists(INDEX_JS) and os.path.exists(CANVAS_JS)


class RecordingCtx:
    """CanvasRenderingContext2D stub: composes the CTM exactly as Canvas2D defines translate / rotate / scale
    (post-multiplication) and records every drawImage with the CTM in force.  No pixels."""

    def __init__(self):
        self.m = [1.0, 0.0, 0.0, 1.0, 0.0, 0.0]   # a b c d e f : x' = a x + c y + e, y' = b x + d y + f
        self.stack = []
        self.draws = []
        self.fills = []
        self.props = {}

    def _mul(self, n):
        a, b, c, d, e, f = self.m
        na, nb, nc, nd, ne, nf = n
        self.m = [a * na + c * nb, b * na + d * nb, a * nc + c * nd, b * nc + d * nd, a * ne + c * nf + e, b * ne + d * nf + f]

    def js_get(self, key):
        fn = {
            "save": lambda t, a: self.stack.append(list(self.m)),
            "restore": lambda t, a: setattr(self, "m", self.stack.pop()),
            "translate": lambda t, a: self._mul([1.0, 0.0, 0.0, 1.0, a[0], a[1]]),
            "scale": lambda t, a: self._mul([a[0], 0.0, 0.0, a[1], 0.0, 0.0]),
            "rotate": lambda t, a: self._mul([math.cos(a[0]), math.sin(a[0]), -math.sin(a[0]), math.cos(a[0]), 0.0, 0.0]),
            "drawImage": lambda t, a: self.draws.append({"ctm": list(self.m), "args": list(a[1:])}),
            "fillRect": lambda t, a: self.fills.append({"ctm": list(self.m), "args": list(a), "fillStyle": self.props.get("fillStyle")}),
            "getImageData": lambda t, a: {},
        }.get(key)
        if fn is not None:
            def wrap(t, a, fn=fn):
                fn(t, a)
                return UNDEF
            return NativeFn(wrap, key)
        return self.props.get(key, UNDEF)

    def js_set(self, key, value):
        self.props[key] = value


def _page_program():
    """The reference's lines, in execution order, with the smallest possible glue (marked GLUE) where the slices
    sit inside control flow that belongs to out-of-scope code (image probing, decoding)."""
    P = INDEX_JS
    parts = [
        ref_lines(P, 1193, 1193),                                 # setData({ isStitching: true, stitchProgress: 1, ... })
        ref_lines(P, 1211, 1213),                                 # totalBytes, bigTask, noCacheInThisStitch
        "let doneProbe = 0;",                                     # GLUE = P:1224
        "for (let idx = 0; idx < images.length; idx++) {",        # GLUE = P:1225 (every record is already prepared:
        "  let current = images[idx];",                           #        P:1226; the else-branch P:1242-1245 runs)
        ref_lines(P, 1243, 1244),                                 # normalisation (the `else` body, P:1242-1245)
        ref_lines(P, 1246, 1248),                                 # probe progress
        "}",                                                      # GLUE = P:1249
        ref_lines(P, 1251, 1386),                                 # layout: base size ... supersample guard loop
        ref_lines(P, 1423, 1433),                                 # white fill, ctx.scale(ss, ss), scaledGap, cursors
        "for (let idx = 0; idx < images.length; idx++) {",        # GLUE = P:1439
        "  const img = images[idx];",                             # GLUE = P:1440
        "  const bmp = bitmaps[idx];",                            # GLUE: the decoded bitmap (P:1441-1520 is the decoder)
        ref_lines(P, 1522, 1557),                                 # naturalW/H, dest rect, drawWithOrientation, cursor, progress
        "}",                                                      # GLUE = P:1572
        ref_lines(P, 1581, 1581),                                 # setData({ stitchProgress: 96 }) after the export
        "const tempPath = { tempFilePath: '' }; let dataDirty = false;",   # GLUE: the exported file (P:1577-1580 is the encoder), P:1208
        ref_lines(P, 1611, 1615),                                 # finalSet ... stitchProgress: 100; this.setData(finalSet)
        # GLUE: hand the locals back to the harness
        "__out.bigTask = bigTask; __out.outW = outW; __out.outH = outH; __out.scaleDown = scaleDown;",
        "__out.MAX_CANVAS_SIDE = MAX_CANVAS_SIDE; __out.MAX_TOTAL_PIXELS = MAX_TOTAL_PIXELS;",
        "__out.targetW = targetW; __out.targetH = targetH; __out.superSample = superSample;",
        "__out.canvasOutW = canvasOutW; __out.canvasOutH = canvasOutH; __out.scaledGap = scaledGap; __out.mode = mode;",
    ]
    return "\n".join(parts)


_PROGRAM_AST = None
_DRAW_AST = None


def run_reference_layout(images, direction, gap, vertical_mode="min", horizontal_mode="min", platform="ios",
                         device_max_size=0, device_max_pixels=0, bitmaps=None):
    """Runs the reference's own layout + placement lines.  `images`: list of dicts with the page's record fields
    (naturalWidth, naturalHeight, width, height, orientation, fileSize; missing = undefined).  Returns a dict with
    the reference's locals, the setData progress values, the pre-clamp integer size, and one entry per
    drawWithOrientation call: the user-space rect it was called with, the CTM at ctx.drawImage and its arguments.
    Raises JSError with the reference's message where the reference throws."""
    global _PROGRAM_AST, _DRAW_AST
    if _PROGRAM_AST is None:
        _PROGRAM_AST = Parser(_page_program()).program()
        _DRAW_AST = Parser(ref_lines(CANVAS_JS, 153, 202)).program()   # function drawWithOrientation(...) { ... }
    it = Interp()

    def num(v):
        return float(v) if isinstance(v, (int, float)) and not isinstance(v, bool) else v

    imgs = [{k: num(v) for k, v in im.items()} for im in images]
    bmps = bitmaps or [{"width": num(im.get("naturalWidth") or im.get("width") or 1),
                        "height": num(im.get("naturalHeight") or im.get("height") or 1)} for im in images]
    bmps = [{k: num(v) for k, v in b.items()} for b in bmps]
    progress, pre = [], {}
    octx = RecordingCtx()
    calls = []

    def set_data(this, a):
        if isinstance(a[0], dict) and "stitchProgress" in a[0]:
            progress.append(a[0]["stitchProgress"])
        return UNDEF

    page = {
        "data": {"verticalStitchMode": vertical_mode, "horizontalStitchMode": horizontal_mode},
        "deviceMaxCanvasSize": float(device_max_size) if device_max_size else UNDEF,
        "deviceMaxCanvasPixels": float(device_max_pixels) if device_max_pixels else UNDEF,
        "setData": NativeFn(set_data, "setData"),
    }
    env = Env(it.globals)
    env.declare("wx", {"getSystemInfoSync": NativeFn(lambda t, a: {"platform": platform})})
    env.declare("images", imgs)
    env.declare("bitmaps", bmps)
    env.declare("direction", direction)
    env.declare("gap", num(gap) if gap is not None else UNDEF)
    env.declare("octx", octx)
    env.declare("__out", {})
    # console.log('计算后尺寸', { outW, outH, mode }) at P:1321 is the one place the pre-clamp integers are visible
    def log(this, a):
        if a and a[0] == "计算后尺寸" and isinstance(a[1], dict):
            pre["preW"], pre["preH"] = a[1]["outW"], a[1]["outH"]
        return UNDEF
    env.declare("console", {"log": NativeFn(log), "warn": NativeFn(lambda t, a: UNDEF), "error": NativeFn(lambda t, a: UNDEF)})
    # utils/canvas.js:153-202 defines drawWithOrientation; the page's calls go through a recorder that notes the
    # arguments (the user-space rect) and then runs the reference's function body unchanged
    it.exec_top(_DRAW_AST, env, UNDEF)
    ref_draw = env.get("drawWithOrientation")

    def draw_recorder(this, a):
        calls.append({"rect": list(a[6:10]), "src": list(a[2:6]), "orientation": a[10] if len(a) > 10 else UNDEF})
        return it.call(ref_draw, UNDEF, a)
    env.set("drawWithOrientation", NativeFn(draw_recorder, "drawWithOrientation"))
    it.exec_top(_PROGRAM_AST, env, page)
    out = env.get("__out")
    res = {k: out[k] for k in out}
    res.update(pre)
    res["progress"] = progress
    res["draws"] = octx.draws
    res["calls"] = calls
    res["fills"] = octx.fills
    res["images"] = imgs
    return res


def run_reference_draw(rect, orientation, super_sample=1.0, bmp_w=8.0, bmp_h=6.0):
    """Runs utils/canvas.js:153-202 alone for one call: returns the CTM at ctx.drawImage and its dest arguments."""
    it = Interp()
    env = Env(it.globals)
    it.exec_top(Parser(ref_lines(CANVAS_JS, 153, 202)).program(), env, UNDEF)
    ctx = RecordingCtx()
    if super_sample != 1.0:
        ctx._mul([float(super_sample), 0.0, 0.0, float(super_sample), 0.0, 0.0])   # P:1426-1428
    dx, dy, dw, dh = (float(v) for v in rect)
    o = float(orientation) if isinstance(orientation, (int, float)) and not isinstance(orientation, bool) else orientation
    it.call(env.get("drawWithOrientation"), UNDEF, [ctx, {"width": bmp_w, "height": bmp_h}, 0.0, 0.0, float(bmp_w), float(bmp_h),
                                                    dx, dy, dw, dh, o])
    assert len(ctx.draws) == 1 and not ctx.stack
    return ctx.draws[0]
