"""Runs the reference's OWN physics source text without a JavaScript engine.

TEST INFRASTRUCTURE ONLY (same rule as the rest of oracle/): importers are tests/ and
tests/golden/make_reference_golden.py.  Nothing of the reference is copied into this
repository: `/root/reference/index.html` is READ at run time, the `CONFIG` literal
(idx:52-75) and the `Body` and `PhysicsSystem` classes (idx:208-444) are sliced out of
the page's module script, parsed by the small JavaScript-subset parser below and
re-emitted, statement by statement with every sub-expression parenthesised as parsed, as
Python source that is then executed.  Python floats are IEEE-754 binary64 like JS Numbers
and `+ - * /` round identically; `Math.pow`/`Math.sqrt` go to the C library (as in
oracle/nbody_oracle.c: the one place V8 may differ by an ulp).

Why: no Node/V8 (or any JS engine) exists in the build image or on the GPU box, so the
Node harness (n-body-sim_b200/addon/run_reference.js) cannot run.  This module is the same
harness with the engine replaced by a transliterator: the statements that execute are the
reference's, not a restatement written from reading them.  It pins oracle/nbody_oracle.c
oracle/collide_oracle.c and oracle/visual_oracle.c's ribbon loop (tests/test_reference_source.py,
tests/test_shim_source.py) and generates the committed vectors tests/golden/ref_js_*.npz.

Supplied from outside the reference, exactly as run_reference.js does:
  * `THREE.Vector3` (three@0.160.0, CDN import idx:24, not on disk): the methods the
    classes call, per the public r160 API;
  * `visuals` (idx:358, 384, 408, 441): call counters, no physics effect;
  * `Math.random` (unseeded in the page): the sequential splitmix64 stream of
    collide_oracle.c, so both sides draw the same numbers;
  * `Body.id` = `Math.random().toString(36).substr(2, 9)` (idx:210): the draw itself stands
    for the string (it is only ever an identity: sorted and joined into a map key, idx:311).

What is executed this way: `CONFIG` (idx:52-75), `Body` and `PhysicsSystem` (idx:208-444), `resetScenario` (idx:744-764),
`VisualSystem.update` (idx:500-597, class `ReferenceVisuals`), and this repository's own `addon/physics_shim.js` installed
on the page's `physics` object (`install_shim`).  The planet fragment shader is oracle/glsl_reference.py's.

The subset: classes and methods with default parameters, function declarations and expressions (closures rebind the
enclosing function's variables through `nonlocal`; a function using `this` must be assigned to an object member and is
bound to that object), expression- and block-bodied arrow functions, const/let, if/else, `for (let i = a; i < b; i++)`
(a `range` when the bound is a name or literal the body leaves alone, else re-evaluated every pass), for-of,
continue/break/return, assignment and compound assignment statements, `++`/`--` statements, `?:`, `|| && !`,
comparisons, arithmetic (`/` with IEEE division by zero), member/index/call chains, `new`, array literals, object
literals with methods, getters and shorthand properties (missing members read as undefined), typed arrays.  Anything
else raises SyntaxError naming the token, so a changed reference fails loudly instead of being half-translated.
"""
from __future__ import annotations

import keyword
import math
import os
import re

REFERENCE_HTML = "/root/reference/index.html"

# ---------------------------------------------------------------- tokenizer

_TOKEN = re.compile(r"""
    (?P<ws>\s+|//[^\n]*|/\*.*?\*/)
  | (?P<num>(?:\d+\.\d*|\.\d+|\d+)(?:[eE][+-]?\d+)?)
  | (?P<str>'(?:[^'\\]|\\.)*'|"(?:[^"\\]|\\.)*")
  | (?P<id>[A-Za-z_$][A-Za-z0-9_$]*)
  | (?P<op>===|!==|=>|==|!=|<=|>=|&&|\|\||\+\+|--|\+=|-=|\*=|/=|[{}()\[\];,.<>+\-*/%!=?:])
""", re.X | re.S)


def tokenize(src: str):
    out, pos = [], 0
    while pos < len(src):
        m = _TOKEN.match(src, pos)
        if not m:
            raise SyntaxError(f"js subset: cannot tokenize at {src[pos:pos + 30]!r}")
        pos = m.end()
        kind = m.lastgroup
        if kind != "ws":
            out.append((kind, m.group(kind)))
    out.append(("eof", ""))
    return out


# ---------------------------------------------------------------- parser (AST = nested tuples)

_ASSIGN_OPS = ("=", "+=", "-=", "*=", "/=")
_BINARY_LEVELS = (("||",), ("&&",), ("===", "!==", "==", "!="), ("<", ">", "<=", ">="), ("+", "-"), ("*", "/", "%"))


class Parser:
    def __init__(self, src: str):
        self.t = tokenize(src)
        self.i = 0

    # token helpers
    def peek(self, k=0):
        return self.t[self.i + k]

    def at(self, text, k=0):
        kind, val = self.t[self.i + k]
        return val == text and kind in ("op", "id")

    def eat(self, text=None, kind=None):
        k, v = self.t[self.i]
        if (text is not None and v != text) or (kind is not None and k != kind):
            raise SyntaxError(f"js subset: expected {text or kind!r}, found {v!r} (token {self.i})")
        self.i += 1
        return v

    def accept(self, text):
        if self.at(text):
            self.i += 1
            return True
        return False

    # program
    def program(self):
        body = []
        while self.peek()[0] != "eof":
            body.append(self.class_decl() if self.at("class") else self.statement())
        return body

    def class_decl(self):
        self.eat("class")
        name = self.eat(kind="id")
        self.eat("{")
        methods = []
        while not self.accept("}"):
            mname = self.eat(kind="id")
            params = self.params()
            methods.append((mname, params, self.block()))
        return ("class", name, methods)

    def params(self):
        self.eat("(")
        ps = []
        while not self.accept(")"):
            pname = self.eat(kind="id")
            default = self.assign() if self.accept("=") else None
            ps.append((pname, default))
            if not self.at(")"):
                self.eat(",")
        return ps

    def block(self):
        self.eat("{")
        body = []
        while not self.accept("}"):
            body.append(self.statement())
        return ("block", body)

    def statement(self):
        if self.at("{"):
            return self.block()
        if self.at("function") and self.peek(1)[0] == "id":
            self.i += 1
            name = self.eat(kind="id")
            params = self.params()
            return ("func", name, params, self.block())
        if self.at("const") or self.at("let") or self.at("var"):
            self.i += 1
            decls = []
            while True:
                name = self.eat(kind="id")
                decls.append((name, self.assign() if self.accept("=") else None))
                if not self.accept(","):
                    break
            self.eat(";")
            return ("decl", decls)
        if self.accept("if"):
            self.eat("(")
            cond = self.expr()
            self.eat(")")
            then = self.statement()
            other = self.statement() if self.accept("else") else None
            return ("if", cond, then, other)
        if self.accept("for"):
            return self.for_statement()
        if self.accept("return"):
            value = None if self.at(";") else self.expr()
            self.eat(";")
            return ("return", value)
        if self.accept("continue"):
            self.eat(";")
            return ("continue",)
        if self.accept("break"):
            self.eat(";")
            return ("break",)
        e = self.expr()
        self.eat(";")
        return ("expr", e)

    def for_statement(self):
        self.eat("(")
        if not (self.at("const") or self.at("let")):
            raise SyntaxError("js subset: for loops must declare their variable")
        self.i += 1
        var = self.eat(kind="id")
        if self.accept("of"):
            it = self.expr()
            self.eat(")")
            return ("forof", var, it, self.statement())
        # for (let v = a; v < b; v++): the only counted form the reference uses (idx:266, 272-273, 422, 751)
        self.eat("=")
        start = self.expr()
        self.eat(";")
        if self.eat(kind="id") != var:
            raise SyntaxError("js subset: for condition must test the loop variable")
        self.eat("<")
        bound = self.expr()
        self.eat(";")
        if self.eat(kind="id") != var:
            raise SyntaxError("js subset: for update must step the loop variable")
        self.eat("++")
        self.eat(")")
        body = self.statement()
        if _assigns(body, var):
            raise SyntaxError("js subset: counted for must not modify its counter")
        if bound[0] == "num" or (bound[0] == "id" and not _assigns(body, bound[1])):
            return ("forrange", var, start, bound, body)
        # any other bound (idx:519 `trailLen - 1`, idx:567 `body.history.length`) is re-evaluated every pass, as in JS
        if _own_continue(body):
            raise SyntaxError("js subset: `continue` in a counted for with a computed bound is not supported")
        return ("forwhile", var, start, bound, body)

    # expressions
    def expr(self):
        return self.assign()

    def assign(self):
        left = self.ternary()
        k, v = self.peek()
        if k == "op" and v in _ASSIGN_OPS:
            self.i += 1
            if left[0] not in ("id", "member", "index"):
                raise SyntaxError("js subset: bad assignment target")
            return ("assign", v, left, self.assign())
        return left

    def ternary(self):
        c = self.binary(0)
        if self.accept("?"):
            a = self.assign()
            self.eat(":")
            return ("cond", c, a, self.assign())
        return c

    def binary(self, level):
        if level == len(_BINARY_LEVELS):
            return self.unary()
        left = self.binary(level + 1)
        while self.peek()[0] == "op" and self.peek()[1] in _BINARY_LEVELS[level]:
            op = self.eat()
            left = ("bin", op, left, self.binary(level + 1))   # left-associative, as in JS
        return left

    def unary(self):
        k, v = self.peek()
        if k == "op" and v in ("!", "-", "+"):
            self.i += 1
            return ("un", v, self.unary())
        return self.postfix()

    def postfix(self):
        e = self.call()
        if self.at("++") or self.at("--"):
            return ("update", self.eat(), e)
        return e

    def call(self):
        e = self.primary()
        while True:
            if self.accept("."):
                e = ("member", e, self.eat(kind="id"))
            elif self.accept("["):
                idx = self.expr()
                self.eat("]")
                e = ("index", e, idx)
            elif self.at("("):
                e = ("call", e, self.args())
            else:
                return e

    def args(self):
        self.eat("(")
        out = []
        while not self.accept(")"):
            out.append(self.assign())
            if not self.at(")"):
                self.eat(",")
        return out

    def _arrow_ahead(self):
        """At '(': is this `(a, b) =>`?"""
        j, depth = self.i, 0
        while True:
            k, v = self.t[j]
            if k == "eof":
                return False
            if k == "op" and v == "(":
                depth += 1
            elif k == "op" and v == ")":
                depth -= 1
                if depth == 0:
                    return self.t[j + 1] == ("op", "=>")
            j += 1

    def _arrow_body(self, ps):
        if not self.at("{"):
            return ("arrow", ps, self.assign())
        return ("arrowblock", ps, self.block())

    def primary(self):
        k, v = self.peek()
        if k == "num":
            self.i += 1
            return ("num", v)
        if k == "str":
            self.i += 1
            return ("str", v[1:-1])
        if k == "id":
            if v == "new":
                self.i += 1
                callee = ("id", self.eat(kind="id"))
                while self.accept("."):
                    callee = ("member", callee, self.eat(kind="id"))
                return ("new", callee, self.args())
            if v == "function":
                self.i += 1
                name = self.eat(kind="id") if self.peek()[0] == "id" else None
                params = self.params()
                return ("funcexpr", name, params, self.block())
            self.i += 1
            if self.at("=>"):
                self.i += 1
                return self._arrow_body([v])
            return ("id", v)
        if v == "(":
            if self._arrow_ahead():
                ps = [p for p, d in self.params()]
                self.eat("=>")
                return self._arrow_body(ps)
            self.i += 1
            e = self.expr()
            self.eat(")")
            return e
        if v == "[":
            self.i += 1
            items = []
            while not self.accept("]"):
                items.append(self.assign())
                if not self.at("]"):
                    self.eat(",")
            return ("array", items)
        if v == "{":
            self.i += 1
            props = []
            while not self.accept("}"):
                kk, key = self.peek()
                if kk not in ("id", "str"):
                    raise SyntaxError(f"js subset: bad property key {key!r}")
                self.i += 1
                if kk == "id" and key == "get" and self.peek()[0] == "id":      # { get name() { ... } }
                    gname = self.eat(kind="id")
                    params = self.params()
                    props.append((gname, ("getter", ("funcexpr", gname, params, self.block()))))
                elif kk == "id" and self.at("("):                              # { name(params) { ... } }
                    params = self.params()
                    props.append((key, ("funcexpr", key, params, self.block())))
                elif kk == "id" and (self.at(",") or self.at("}")):                # { name }  =  { name: name }
                    props.append((key, ("id", key)))
                else:
                    self.eat(":")
                    props.append((key[1:-1] if kk == "str" else key, self.assign()))
                if not self.at("}"):
                    self.eat(",")
            return ("object", props)
        raise SyntaxError(f"js subset: unexpected token {v!r} (token {self.i})")


def _assigns(node, name) -> bool:
    """Does the statement/expression tree assign to (or ++/--) the plain identifier `name`?"""
    if isinstance(node, tuple):
        if node and node[0] == "assign" and node[2] == ("id", name):
            return True
        if node and node[0] == "update" and node[2] == ("id", name):
            return True
        if node and node[0] == "decl" and any(n == name for n, _ in node[1]):
            return True
        return any(_assigns(c, name) for c in node[1:])
    if isinstance(node, list):
        return any(_assigns(c, name) for c in node)
    return False


def _walk(node):
    if isinstance(node, tuple):
        yield node
        for c in node[1:]:
            yield from _walk(c)
    elif isinstance(node, list):
        for c in node:
            yield from _walk(c)


_FUNCTION_NODES = ("func", "funcexpr", "arrowblock", "arrow")


def _walk_shallow(node, top=True):
    """Like _walk, but does not enter nested functions (their names are their own)."""
    if isinstance(node, tuple):
        if not top and node and node[0] in _FUNCTION_NODES:
            if node[0] == "func":
                yield ("decl", [(node[1], None)])      # a nested function declaration binds its name here
            return
        yield node
        for c in node[1:]:
            yield from _walk_shallow(c, False)
    elif isinstance(node, list):
        for c in node:
            yield from _walk_shallow(c, False)


def _scope_names(params, body):
    """(declared, assigned) plain identifiers of one function body, nested functions excluded."""
    declared, assigned = {p if isinstance(p, str) else p[0] for p in params}, set()
    for n in _walk_shallow(body, False):
        if not n:
            continue
        if n[0] == "decl":
            declared |= {name for name, _ in n[1]}
        elif n[0] in ("forrange", "forwhile", "forof"):
            declared.add(n[1])
        elif n[0] in ("assign", "update") and isinstance(n[2], tuple) and n[2][0] == "id":
            assigned.add(n[2][1])
    return declared, assigned


def _uses_this(body) -> bool:
    return any(n == ("id", "this") for n in _walk_shallow(body, False))


def _own_continue(node) -> bool:
    """A `continue` that belongs to the loop whose body `node` is (not to a loop nested inside it)."""
    if isinstance(node, tuple):
        if node and node[0] == "continue":
            return True
        if node and node[0] in ("forrange", "forwhile", "forof", "arrowblock"):
            return False
        return any(_own_continue(c) for c in node[1:])
    if isinstance(node, list):
        return any(_own_continue(c) for c in node)
    return False


# ---------------------------------------------------------------- Python emitter

_CONSTANTS = {"true": "True", "false": "False", "null": "None", "undefined": "None"}
_OPS = {"===": "==", "!==": "!=", "||": "or", "&&": "and"}


def _name(n: str) -> str:
    if n in _CONSTANTS:
        return _CONSTANTS[n]
    n = n.replace("$", "_S_")
    return n + "_" if keyword.iskeyword(n) else n


_HOIST = []      # function-valued expressions met while emitting a statement's expressions: (name, params, block, takes_self)
_COUNTER = [0]
_SCOPES = []     # declared names of the enclosing functions, innermost last
_THIS_FN_OK = [False]


def _emit_function(name, params, block, ind, out, takes_self=False):
    """def name(params): body -- with `nonlocal` for the enclosing functions' variables it assigns (JS closures rebind freely)."""
    ps = (["self"] if takes_self else []) + [_name(p) if isinstance(p, str) else _name(p[0]) + (f"={emit_expr(p[1])}" if p[1] is not None else "")
                                             for p in params]
    declared, assigned = _scope_names(params, block)
    outer = sorted(x for x in assigned - declared if any(x in sc for sc in _SCOPES))
    free = sorted(x for x in assigned - declared if not any(x in sc for sc in _SCOPES))
    if free:
        raise SyntaxError(f"js subset: function {name} assigns to undeclared or module-level name(s) {free}")
    out.append(f"{'    ' * ind}def {name}({', '.join(ps)}):")
    if outer:
        out.append(f"{'    ' * (ind + 1)}nonlocal {', '.join(_name(x) for x in outer)}")
    _SCOPES.append(declared)
    try:
        emit_body(block, ind + 1, out)
    finally:
        _SCOPES.pop()


def _flush_hoisted(ind: int, out: list):
    pending = _HOIST[:]          # (functions hoisted while THESE bodies are emitted flush inside them, at their own statements)
    del _HOIST[:]
    for name, params, block, takes_self in pending:
        _emit_function(name, params, block, ind, out, takes_self)


def emit_expr(e) -> str:
    k = e[0]
    if k == "arrowblock":
        _COUNTER[0] += 1
        name = f"_arrow_{_COUNTER[0]}"
        _HOIST.append((name, e[1], e[2], False))
        return name
    if k == "funcexpr":      # `this` inside is the object the function is assigned to: bound where it is assigned (emit_stmt)
        if _uses_this(e[3]) and not _THIS_FN_OK[0]:
            raise SyntaxError("js subset: a function expression using `this` is only supported as `obj.member = function ...`")
        _COUNTER[0] += 1
        name = f"_fn_{_COUNTER[0]}_{_name(e[1] or 'anonymous')}"
        _HOIST.append((name, e[2], e[3], _uses_this(e[3])))
        return name
    if k == "getter":
        return f"JSGetter({emit_expr(e[1])})"
    if k == "num":
        return e[1]
    if k == "str":
        return repr(e[1])
    if k == "id":
        return "self" if e[1] == "this" else _name(e[1])
    if k == "member":
        return f"{emit_expr(e[1])}.{_name(e[2])}"
    if k == "index":
        return f"{emit_expr(e[1])}[{emit_expr(e[2])}]"
    if k in ("call", "new"):
        return f"{emit_expr(e[1])}({', '.join(emit_expr(a) for a in e[2])})"
    if k == "bin":
        a, b = emit_expr(e[2]), emit_expr(e[3])
        if e[1] == "/":
            return f"_div({a}, {b})"
        return f"({a} {_OPS.get(e[1], e[1])} {b})"
    if k == "un":
        return f"(not {emit_expr(e[2])})" if e[1] == "!" else f"({e[1]}{emit_expr(e[2])})"
    if k == "cond":
        return f"({emit_expr(e[2])} if {emit_expr(e[1])} else {emit_expr(e[3])})"
    if k == "arrow":
        return f"(lambda {', '.join(_name(p) for p in e[1])}: {emit_expr(e[2])})"
    if k == "array":
        return f"JSArray([{', '.join(emit_expr(a) for a in e[1])}])"
    if k == "object":
        return "JSObject({" + ", ".join(f"{key!r}: {emit_expr(v)}" for key, v in e[1]) + "})"
    raise SyntaxError(f"js subset: {k} is only supported as a statement")


def emit_stmt(s, ind: int, out: list):
    pad = "    " * ind
    k = s[0]

    def line(text):          # the statement's own expressions are emitted first, so hoisted defs land above it
        _flush_hoisted(ind, out)
        out.append(pad + text)

    if k == "block":
        if not s[1]:
            out.append(pad + "pass")
        for c in s[1]:
            emit_stmt(c, ind, out)
    elif k == "func":
        _flush_hoisted(ind, out)
        _emit_function(_name(s[1]), s[2], s[3], ind, out)
    elif k == "decl":
        for name, value in s[1]:
            line(f"{_name(name)} = {emit_expr(value) if value is not None else 'None'}")
    elif k == "if":
        line(f"if {emit_expr(s[1])}:")
        emit_body(s[2], ind + 1, out)
        if s[3] is not None:
            out.append(pad + "else:")
            emit_body(s[3], ind + 1, out)
    elif k == "forof":
        line(f"for {_name(s[1])} in {emit_expr(s[2])}:")
        emit_body(s[3], ind + 1, out)
    elif k == "forrange":
        line(f"for {_name(s[1])} in range({emit_expr(s[2])}, {emit_expr(s[3])}):")
        emit_body(s[4], ind + 1, out)
    elif k == "forwhile":
        line(f"{_name(s[1])} = {emit_expr(s[2])}")
        line(f"while ({_name(s[1])} < {emit_expr(s[3])}):")
        emit_body(s[4], ind + 1, out)
        out.append(f"{pad}    {_name(s[1])} += 1")
    elif k == "return":
        line(f"return {emit_expr(s[1]) if s[1] is not None else ''}".rstrip())
    elif k in ("continue", "break"):
        out.append(pad + k)
    elif k == "expr":
        e = s[1]
        if e[0] == "assign" and e[3][0] == "funcexpr" and _uses_this(e[3][3]):
            if e[1] != "=" or e[2][0] != "member":
                raise SyntaxError("js subset: a function using `this` must be assigned to an object member")
            _THIS_FN_OK[0] = True
            try:
                fn = emit_expr(e[3])
            finally:
                _THIS_FN_OK[0] = False
            line(f"{emit_expr(e[2])} = _bind({fn}, {emit_expr(e[2][1])})")
        elif e[0] == "assign":
            if e[1] == "/=":
                line(f"{emit_expr(e[2])} = _div({emit_expr(e[2])}, {emit_expr(e[3])})")
            else:
                line(f"{emit_expr(e[2])} {e[1]} {emit_expr(e[3])}")
        elif e[0] == "update":
            line(f"{emit_expr(e[2])} {'+=' if e[1] == '++' else '-='} 1")
        else:
            line(emit_expr(e))
    else:
        raise SyntaxError(f"js subset: cannot emit {k}")


def emit_body(s, ind, out):
    n = len(out)
    emit_stmt(s, ind, out)
    if len(out) == n:
        out.append("    " * ind + "pass")


def emit_program(prog) -> str:
    out = []
    for node in prog:
        if node[0] == "class":
            out.append(f"class {_name(node[1])}:")
            for mname, params, body in node[2]:
                _emit_function("__init__" if mname == "constructor" else _name(mname), params, body, 1, out, takes_self=True)
            out.append("")
        else:
            emit_stmt(node, 0, out)
    return "\n".join(out) + "\n"


def transliterate(js_src: str) -> str:
    return emit_program(Parser(js_src).program())


# ---------------------------------------------------------------- runtime the emitted code runs against

def _div(a, b):
    try:
        return a / b
    except ZeroDivisionError:                # JS: x/0 = +-Infinity, 0/0 = NaN
        a, b = float(a), float(b)
        if a != a or a == 0.0:
            return math.nan
        return math.inf if (a > 0) == (math.copysign(1.0, b) > 0) else -math.inf


class JSArray(list):
    @property
    def length(self):
        return len(self)

    def push(self, *items):
        self.extend(items)
        return len(self)

    def shift(self):
        return self.pop(0) if self else None

    def filter(self, fn):
        return JSArray(x for i, x in enumerate(self) if fn(x, i))

    def forEach(self, fn):
        i = 0
        while i < len(self):                 # live length, like Array.prototype.forEach over a dense array
            fn(self[i])
            i += 1

    def map(self, fn):
        return JSArray(fn(x) for x in self)

    def unshift(self, *items):
        self[0:0] = items
        return len(self)

    def sort(self, cmp=None):                # ids are numbers here (see module docstring): numeric order
        assert cmp is None
        list.sort(self)
        return self

    def join(self, sep=","):                 # only ever a Map key (idx:311): a hashable of the sorted ids
        return tuple(self)


class JSGetter:
    def __init__(self, fn):
        self.fn = fn


class JSObject:
    def __init__(self, d):
        getters = {k: v.fn for k, v in d.items() if isinstance(v, JSGetter)}
        self.__dict__.update({k: v for k, v in d.items() if k not in getters})
        self.__dict__["_getters"] = getters

    def __getattr__(self, key):              # only reached for names not in __dict__: a getter, else `undefined`
        getters = self.__dict__.get("_getters", {})
        return getters[key]() if key in getters else None

    def __getitem__(self, key):
        return self.__dict__[key]

    def __setitem__(self, key, value):
        self.__dict__[key] = value


class _ObjectNS:
    """`Object.defineProperty(obj, key, descriptor)` for the shim's typed-array-backed Body views (addon/physics_shim.js):
    an accessor descriptor ({get, set}) or a data descriptor ({value}) on ONE object.  Python keeps accessors on classes, so
    the object moves to a private subclass of its class that carries the properties (slots classes included)."""

    @staticmethod
    def defineProperty(obj, key, desc):
        cls = type(obj)
        if not cls.__dict__.get("_js_own_class", False):
            body = {"_js_own_class": True}
            if not hasattr(obj, "__dict__"):
                body["__slots__"] = ()
            sub = type(cls.__name__, (cls,), body)
            obj.__class__ = sub
            cls = sub
        get, set_ = desc.get, desc.set
        if get is not None or set_ is not None:
            def fget(self, _g=get):
                return _g()

            def fset(self, value, _s=set_):
                if _s is None:
                    raise TypeError(f"js: property {key!r} has no setter")
                _s(value)
            setattr(cls, key, property(fget, fset))
        else:
            if key in cls.__dict__:
                delattr(cls, key)                # the accessor goes; the base class's slot / the instance dict takes over
            object.__setattr__(obj, key, desc.value)
        return obj


def _bind(fn, obj):
    """`obj.name = function () { ... this ... }`: calls through obj see it as `this`."""
    import types
    return types.MethodType(fn, obj)


def typed_array(dtype):
    """Float64Array / Float32Array / Uint32Array: numpy storage (stores round like the JS typed array), `length`, `subarray`."""
    import numpy as np

    class TypedArray(np.ndarray):
        def __new__(cls, n):
            return np.zeros(int(n), dtype=dtype).view(cls)

        @property
        def length(self):
            return self.shape[0]

        def subarray(self, begin, end=None):
            return self[int(begin):None if end is None else int(end)]

        def __getitem__(self, key):          # an element read is a JS Number (binary64), whatever the storage type
            v = np.ndarray.__getitem__(self, key)
            return v.item() if isinstance(v, np.generic) else v

    TypedArray.__name__ = np.dtype(dtype).name
    return TypedArray


class JSMap:
    def __init__(self):
        self._d = {}

    def get(self, k):
        return self._d.get(k)                # undefined -> None (falsy, so `|| 0` works, idx:314)

    def set(self, k, v):
        self._d[k] = v
        return self

    def has(self, k):
        return k in self._d

    def delete(self, k):
        return self._d.pop(k, None) is not None

    def clear(self):
        self._d.clear()

    def keys(self):
        return list(self._d)                 # JS Map iteration tolerates deletion (idx:375-379): snapshot

    @property
    def size(self):
        return len(self._d)


class JSSet:
    def __init__(self):
        self._s = set()

    def has(self, k):
        return k in self._s

    def add(self, k):
        self._s.add(k)
        return self

    @property
    def size(self):
        return len(self._s)


class Vector3:
    """three@0.160.0 Vector3 (public API; CDN import idx:24): the members idx:208-444 call."""
    __slots__ = ("x", "y", "z")

    def __init__(self, x=0, y=0, z=0):
        self.x, self.y, self.z = x, y, z

    def clone(self):
        return Vector3(self.x, self.y, self.z)

    def copy(self, v):
        self.x, self.y, self.z = v.x, v.y, v.z
        return self

    def set(self, x, y, z):
        self.x, self.y, self.z = x, y, z
        return self

    def add(self, v):
        self.x += v.x
        self.y += v.y
        self.z += v.z
        return self

    def sub(self, v):
        self.x -= v.x
        self.y -= v.y
        self.z -= v.z
        return self

    def addScaledVector(self, v, s):
        self.x += v.x * s
        self.y += v.y * s
        self.z += v.z * s
        return self

    def subVectors(self, a, b):
        self.x = a.x - b.x
        self.y = a.y - b.y
        self.z = a.z - b.z
        return self

    def multiplyScalar(self, s):
        self.x *= s
        self.y *= s
        self.z *= s
        return self

    def divideScalar(self, s):
        return self.multiplyScalar(_div(1, s))

    def dot(self, v):
        return self.x * v.x + self.y * v.y + self.z * v.z

    def lengthSq(self):
        return self.x * self.x + self.y * self.y + self.z * self.z

    def length(self):
        return math.sqrt(self.x * self.x + self.y * self.y + self.z * self.z)

    def normalize(self):
        return self.divideScalar(self.length() or 1)

    def crossVectors(self, a, b):
        ax, ay, az, bx, by, bz = a.x, a.y, a.z, b.x, b.y, b.z
        self.x = ay * bz - az * by
        self.y = az * bx - ax * bz
        self.z = ax * by - ay * bx
        return self

    def setScalar(self, s):
        self.x = self.y = self.z = s
        return self


class _Color:
    def __init__(self, *rgb):
        self.rgb = rgb


class _Draw(float):
    """A Math.random() result.  `.toString(36).substr(2, 9)` (idx:210) yields the draw itself."""

    def toString(self, radix=10):
        return self

    def substr(self, start, length=None):
        return float(self)


class _Math:
    PI = math.pi
    pow = staticmethod(math.pow)
    sqrt = staticmethod(math.sqrt)
    floor = staticmethod(math.floor)
    min = staticmethod(min)
    max = staticmethod(max)
    abs = staticmethod(abs)
    cos = staticmethod(math.cos)
    sin = staticmethod(math.sin)

    def __init__(self, seed):
        self.state = seed & 0xFFFFFFFFFFFFFFFF
        self.draws = 0

    def random(self):                        # collide_oracle.c rnd(): sequential splitmix64, 53 bits
        M = 0xFFFFFFFFFFFFFFFF
        self.state = (self.state + 0x9E3779B97F4A7C15) & M
        z = self.state
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & M
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & M
        z ^= z >> 31
        self.draws += 1
        return _Draw((z >> 11) * (1.0 / 9007199254740992.0))


class _Visuals:
    def __init__(self):
        self.flash = self.explosion = self.gc = 0

    def triggerFlash(self, pos, energy):
        self.flash += 1

    def spawnExplosion(self, pos, temp):
        self.explosion += 1

    def garbageCollect(self, bodies):
        self.gc += 1


class _Three:
    Vector3 = Vector3
    Color = _Color


def _slice(text, start, end):
    a = text.find(start)
    b = text.find(end, a)
    if a < 0 or b < 0:
        raise RuntimeError(f"cannot find {start!r} .. {end!r} in the reference page")
    return text[a:b]


TYPES = ("rock", "metal", "ice")


class Reference:
    """The page's CONFIG, Body and PhysicsSystem, executed from its own source text."""

    def __init__(self, html=REFERENCE_HTML, seed=0x5EED0001, collisions=True):
        with open(html, encoding="utf-8") as f:
            page = f.read()
        config_js = _slice(page, "const CONFIG = {", "const V3")          # idx:52-75
        classes_js = _slice(page, "class Body {", "// 4. VISUALS")         # idx:208-444
        classes_js = classes_js[:classes_js.rstrip().rfind("}") + 1]
        scenario_js = _slice(page, "function resetScenario(type) {", "resetScenario('galaxy');")   # idx:744-764
        self.python_source = transliterate(config_js + "\n" + classes_js + "\n" + scenario_js)
        self.Math, self.visuals = _Math(seed), _Visuals()
        ns = {"THREE": _Three, "Math": self.Math, "visuals": self.visuals, "JSArray": JSArray, "JSObject": JSObject,
              "Map": JSMap, "Set": JSSet, "_div": _div, "V3": Vector3}   # V3 = (x,y,z) => new THREE.Vector3(x,y,z), idx:77
        exec(compile(self.python_source, "<index.html idx:52-75,208-444 transliterated>", "exec"), ns)
        self.CONFIG, self.Body, self.PhysicsSystem = ns["CONFIG"], ns["Body"], ns["PhysicsSystem"]
        if not collisions:   # gravity-only parity runs: idx:256 becomes a no-op, as in run_reference.js
            self.PhysicsSystem.resolveCollisions = lambda self_, dt: None
        self.physics = ns["physics"] = self.PhysicsSystem()                # `const physics = new PhysicsSystem()`, the page's global
        self._reset = ns["resetScenario"]

    def resetScenario(self, type="galaxy"):
        """The page's own scene generator (idx:744-764), drawing from the seeded Math.random."""
        self._reset(type)

    def add(self, mass, pos, vel, type="rock", temp=0.0, acc=None):
        b = self.Body(mass, Vector3(*map(float, pos)), Vector3(*map(float, vel)), type, temp)
        if acc is not None:
            b.acc.set(*map(float, acc))
        return self.physics.addBody(b)

    def integrate(self, dt, nsteps=1):
        for _ in range(nsteps):
            self.physics.integrate(dt)

    def state(self):
        import numpy as np
        bodies = self.physics.bodies
        n = len(bodies)
        st = {"pos_mass": np.zeros((n, 4)), "vel": np.zeros((n, 4)), "acc": np.zeros((n, 4)), "temp": np.zeros(n),
              "radius": np.zeros(n), "type": np.zeros(n, dtype=np.int32), "id": np.zeros(n)}
        for i, b in enumerate(bodies):
            st["pos_mass"][i] = (b.pos.x, b.pos.y, b.pos.z, b.mass)
            st["vel"][i, :3] = (b.vel.x, b.vel.y, b.vel.z)
            st["acc"][i, :3] = (b.acc.x, b.acc.y, b.acc.z)
            st["temp"][i], st["radius"][i], st["id"][i] = b.temperature, b.radius, b.id
            st["type"][i] = TYPES.index(b.type)
        return st


# ---------------------------------------------------------------- the renderer feed: VisualSystem.update (idx:500-597)

class _Bag:
    """A plain JS object / three.js object whose members the page only stores."""

    def __init__(self, *args, **kw):
        self.args = args
        self.__dict__.update(kw)


class _BufferGeometry(_Bag):
    def __init__(self):
        self.attributes = _Bag()
        self.index = None

    def setIndex(self, indices):
        self.index = indices

    def setAttribute(self, name, attr):
        setattr(self.attributes, name, attr)


class _BufferAttribute(_Bag):
    def __init__(self, array, itemSize):
        self.array, self.itemSize, self.needsUpdate = array, itemSize, False


class _Mesh(_Bag):
    def __init__(self, geometry, material):
        self.geometry, self.material = geometry, material
        self.position, self.scale, self.rotation = Vector3(), Vector3(1, 1, 1), Vector3()
        self.frustumCulled = True


class _ThreeVisual(_Three):
    Mesh = _Mesh
    BufferGeometry = _BufferGeometry
    BufferAttribute = _BufferAttribute
    MeshBasicMaterial = _Bag
    AdditiveBlending = 2
    DoubleSide = 2


def _float32_array(n):
    import numpy as np
    return np.zeros(int(n), dtype=np.float32)      # stores round to binary32, like a JS Float32Array


class ReferenceVisuals:
    """`VisualSystem.update(bodies, dt)` executed from the page's source (idx:500-597): per body the
    `history` unshift/pop (idx:540, 564-565) and the camera-facing ribbon vertices (idx:567-592) written
    into a Float32Array.  The method is sliced out on its own: the constructor, materials, particles and
    flashes (WebGL objects, `Math.random`-driven, no effect on the ribbons) are stand-ins, and the scene
    graph objects it creates are inert records."""

    def __init__(self, reference: "Reference", camera=(0.0, 80.0, 150.0), html=REFERENCE_HTML):
        with open(html, encoding="utf-8") as f:
            page = f.read()
        body = _slice(page, "update(bodies, dt) {", "garbageCollect(activeBodies) {")
        self.python_source = transliterate("class VisualUpdate {\n" + body + "\n}")
        self.Math = _Math(0xC0FFEE)                 # its own stream: idx:557 draws only for bodies hotter than 50
        self.clock = 0.0
        ns = {"THREE": _ThreeVisual, "Math": self.Math, "CONFIG": reference.CONFIG, "JSArray": JSArray, "JSObject": _Bag_from_dict,
              "Map": JSMap, "Set": JSSet, "_div": _div, "Float32Array": _float32_array,
              "performance": _Bag(now=lambda: self.clock)}
        exec(compile(self.python_source, "<index.html idx:500-597 transliterated>", "exec"), ns)
        v = self.v = ns["VisualUpdate"]()
        v.scene, v.camera = _Bag(add=lambda obj: None), _Bag(position=Vector3(*map(float, camera)))
        v.meshMap, v.trailMap, v.geo = JSMap(), JSMap(), _Bag()                       # idx:454-458
        v.vecToCam, v.vecDir, v.vecPerp = Vector3(), Vector3(), Vector3()            # idx:480-482
        v.createPlanetMaterial = lambda type: _Bag(uniforms=_Bag(time=_Bag(value=0), temperature=_Bag(value=0)))
        v.spawnTrail = lambda body: None
        v.updateParticles = lambda dt: None
        self.reference = reference

    def update(self, dt):
        self.clock += 1000.0 * dt
        self.v.update(self.reference.physics.bodies, dt)

    def ribbons(self):
        """float32 [n, trailLength, 6] in body order: the `position` arrays the page hands to WebGL."""
        import numpy as np
        return np.stack([self.v.trailMap.get(b.id).positions.reshape(-1, 6) for b in self.reference.physics.bodies])

    def histories(self):
        import numpy as np
        return np.array([[(p.x, p.y, p.z) for p in b.history] for b in self.reference.physics.bodies])


def _Bag_from_dict(d):
    return _Bag(**d)


# ---------------------------------------------------------------- the drop-in: the page with its integrate() replaced by the shim

SHIM_JS = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "n-body-sim_b200", "addon", "physics_shim.js")


def install_shim(reference: "Reference", addon, **opts):
    """Executes n-body-sim_b200/addon/physics_shim.js (this repository's JS, transliterated like the page) against the
    page's own `physics` and `CONFIG`: `installGpuIntegrate(physics, CONFIG, addon, opts)` replaces
    `PhysicsSystem.integrate` (idx:247-262) and `computeGravity` on that one object; Body, addBody, resolveCollisions,
    mergeBodies, fractureBodies stay the page's.  `addon` is any object with the N-API addon's methods
    (create/destroy/upload/download/step/stepHost/kick1Drift/force/kick2/setHeatDecay/setRadius/contacts).
    Returns the shim's handle (`dispose()`, `ctx`)."""
    with open(SHIM_JS, encoding="utf-8") as f:
        src = f.read()
    module = JSObject({})
    ns = {"JSArray": JSArray, "JSObject": JSObject, "JSGetter": JSGetter, "_bind": _bind, "_div": _div, "Math": reference.Math,
          "Float64Array": typed_array("float64"), "Float32Array": typed_array("float32"), "Uint32Array": typed_array("uint32"),
          "Object": _ObjectNS, "module": module}
    exec(compile(transliterate(src), "<physics_shim.js transliterated>", "exec"), ns)
    return module.exports.installGpuIntegrate(reference.physics, reference.CONFIG, addon, JSObject(opts))


def available(html=REFERENCE_HTML) -> bool:
    return os.path.exists(html)
