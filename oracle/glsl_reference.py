"""Runs the reference's OWN planet shader text (GLSL ES: fragment idx:99-202, vertex idx:83-97) without a graphics stack.

TEST INFRASTRUCTURE ONLY.  `/root/reference/index.html` is read at run time, the `planetFragmentShader` template
string is sliced out, parsed by the GLSL-subset parser below (it reuses oracle/js_reference.py's tokenizer and
expression grammar) and re-emitted one statement for one statement as Python over numpy float32 values, then executed
for single fragments.  Every GLSL operation becomes one binary32 operation (numpy float32 scalars and arrays round
after every `+ - * /`, no fused multiply-add), which is the arithmetic oracle/visual_oracle.c and the CUDA renderer
(csrc/render.cu, -fmad=false) commit to.  It pins the C restatement's STRUCTURE -- constants, swizzles, operation
order, which noise lattice point goes where -- against the shader's text (tests/test_reference_source.py).

What GLSL leaves to the implementation is fixed the way visual_oracle.c fixes it, and listed here so the pin is honest:
  dot            products added left to right                       normalize(v)   v / sqrt(dot(v, v))
  mix(a, b, t)   a*(1 - t) + b*t                                    clamp(x,l,h)   min(max(x, l), h)
  smoothstep     t = clamp((x - e0)/(e1 - e0), 0, 1); t*t*(3 - 2*t)  step(e, x)     x < e ? 0 : 1
  pow(x, 3.0)    f = max(x, 0); f*f*f (pow of a negative base is undefined in GLSL; only this exponent occurs)
  literals and constant expressions (1.0/6.0, 1.0/289.0) are binary32
Parity against a real WebGL driver stays unpinned (no graphics stack exists here or on the GPU box).

The subset is what idx:99-202 uses: uniform/varying declarations, function definitions (overloads must have identical
bodies: mod289), `[const] type name = expr;`, assignment and `*=`, `return`, calls, constructors vec2/3/4, read
swizzles, `+ - * /`, unary minus; for the vertex shader mat3/mat4 uniforms whose `*` is the matrix product (class Mat).
Anything else raises SyntaxError.
"""
from __future__ import annotations

import numpy as np

from . import js_reference as jr

F = np.float32
_TYPES = ("float", "vec2", "vec3", "vec4", "mat3", "mat4", "void")


class _Parser(jr.Parser):
    def shader(self):
        funcs, uniforms = [], []
        while self.peek()[0] != "eof":
            if self.at("uniform") or self.at("varying"):
                self.i += 1
                self.eat(kind="id")
                uniforms.append(self.eat(kind="id"))
                self.eat(";")
                continue
            rtype = self.eat(kind="id")
            if rtype not in _TYPES:
                raise SyntaxError(f"glsl subset: unexpected {rtype!r} at top level")
            name = self.eat(kind="id")
            self.eat("(")
            params = []
            while not self.accept(")"):
                if self.eat(kind="id") not in _TYPES:
                    raise SyntaxError("glsl subset: typed parameters expected")
                params.append(self.eat(kind="id"))
                if not self.at(")"):
                    self.eat(",")
            self.eat("{")
            body = []
            while not self.accept("}"):
                body.append(self.glsl_statement())
            funcs.append((name, params, body))
        return funcs, uniforms

    def glsl_statement(self):
        if self.accept("return"):
            e = self.expr()
            self.eat(";")
            return ("return", e)
        self.accept("const")
        if self.peek()[1] in _TYPES and self.peek(1)[0] == "id":      # declaration: type name = expr;
            self.i += 1
            name = self.eat(kind="id")
            self.eat("=")
            e = self.expr()
            self.eat(";")
            return ("set", name, e, True)
        name = self.eat(kind="id")
        op = self.eat()
        if op not in ("=", "*=", "+=", "-="):
            raise SyntaxError(f"glsl subset: unsupported statement at {name} {op}")
        e = self.expr()
        self.eat(";")
        return ("set", name, e if op == "=" else ("bin", op[0], ("id", name), e), False)


def _emit(e) -> str:
    k = e[0]
    if k == "num":
        return f"F({e[1]})"
    if k == "id":
        return e[1]
    if k == "member":
        return f"swz({_emit(e[1])}, {e[2]!r})"
    if k == "call":
        if e[1][0] != "id":
            raise SyntaxError("glsl subset: only plain function calls")
        name = e[1][1]
        args = ", ".join(_emit(a) for a in e[2])
        return f"vec({name[3]}, {args})" if name in ("vec2", "vec3", "vec4") else f"{name}({args})"
    if k == "bin" and e[1] in "+-*/":
        return f"({_emit(e[2])} {e[1]} {_emit(e[3])})"
    if k == "un" and e[1] == "-":
        return f"(-{_emit(e[2])})"
    raise SyntaxError(f"glsl subset: unsupported expression {k} {e[1] if len(e) > 1 else ''}")


def transliterate(glsl: str) -> str:
    funcs, _ = _Parser(glsl).shader()
    out, seen = [], {}
    for name, params, body in funcs:
        lines = [f"    {s[1]} = {_emit(s[2])}" if s[0] == "set" else f"    return {_emit(s[1])}" for s in body]
        if name == "main":       # varyings and gl_* outputs: the names main() assigns without declaring
            outs = sorted({st[1] for st in body if st[0] == "set" and not st[3]} - {st[1] for st in body if st[0] == "set" and st[3]})
            if outs:
                lines.insert(0, "    global " + ", ".join(outs))
        text = "\n".join(lines)
        if name in seen:
            if seen[name] != text:
                raise SyntaxError(f"glsl subset: overloads of {name} differ in more than their types")
            continue
        seen[name] = text
        out.append(f"def {name}({', '.join(params)}):\n{text}\n")
    return "\n".join(out)


# ---- built-ins, binary32, fixed as in visual_oracle.c (module docstring) ----------------------------------------------

def vec(n, *args):
    parts = [np.atleast_1d(np.asarray(a, dtype=F)) for a in args]
    v = np.concatenate(parts)
    n = int(n)
    if v.shape[0] == 1:
        v = np.repeat(v, n)
    if v.shape[0] != n:
        raise ValueError(f"vec{n} from {v.shape[0]} components")
    return v


def swz(v, comps):
    idx = ["xyzw".index(c) for c in comps]
    v = np.asarray(v)
    return v[idx[0]] if len(idx) == 1 else v[idx]


def dot(a, b):
    p = np.asarray(a, dtype=F) * np.asarray(b, dtype=F)
    s = p[0]
    for t in p[1:]:
        s = F(s + t)
    return s


def step(edge, x):
    return np.where(np.asarray(x) < np.asarray(edge), F(0), F(1)).astype(F)[()]


def mix(a, b, t):
    return a * (F(1) - t) + b * t


def clamp(x, lo, hi):
    return np.minimum(np.maximum(x, lo), hi)


def smoothstep(e0, e1, x):
    t = clamp((x - e0) / (e1 - e0), F(0), F(1))
    return t * t * (F(3) - F(2) * t)


def normalize(v):
    return v / np.sqrt(dot(v, v))


def glsl_pow(x, y):
    if y != F(3):
        raise ValueError("glsl subset: pow is fixed for the exponent 3.0 only (idx:184)")
    f = np.maximum(x, F(0))
    return f * f * f


_BUILTINS = {"F": F, "vec": vec, "swz": swz, "dot": dot, "step": step, "mix": mix, "clamp": clamp, "smoothstep": smoothstep,
             "normalize": normalize, "pow": glsl_pow, "floor": np.floor, "abs": np.abs, "min": np.minimum, "max": np.maximum}


class Mat(np.ndarray):
    """mat3 / mat4: `*` with a matrix or a vector is the linear-algebra product, as in GLSL (row-major storage here)."""

    def __new__(cls, rows):
        return np.asarray(rows, dtype=F).view(cls)

    def __mul__(self, other):
        o = np.asarray(other, dtype=F)
        a = np.asarray(self)
        if o.ndim == 0:
            return Mat(a * o)
        out = np.zeros(a.shape[0] if o.ndim == 1 else (a.shape[0], o.shape[1]), dtype=F)
        for k in range(a.shape[1]):                      # products added left to right, one rounding each
            out = (out + (a[:, k] * o[k] if o.ndim == 1 else np.outer(a[:, k], o[k]))).astype(F)
        return Mat(out) if o.ndim == 2 else out


class PlanetVertexShader:
    """`planetVertexShader` of the page (idx:83-97), executable per vertex: the varyings the fragment shader receives."""

    def __init__(self, html=jr.REFERENCE_HTML):
        with open(html, encoding="utf-8") as f:
            page = f.read()
        a = page.index("const planetVertexShader = `") + len("const planetVertexShader = `")
        self.glsl = page[a:page.index("`", a)]
        self.python_source = transliterate(self.glsl)
        self.ns = dict(_BUILTINS)
        exec(compile(self.python_source, "<planetVertexShader idx:83-97 transliterated>", "exec"), self.ns)

    def vertex(self, position, normal, modelMatrix, viewMatrix, projectionMatrix, normalMatrix, uv=(0.0, 0.0)):
        ns = self.ns
        ns.update(position=vec(3, *position), normal=vec(3, *normal), uv=vec(2, *uv), modelMatrix=Mat(modelMatrix), viewMatrix=Mat(viewMatrix),
                  projectionMatrix=Mat(projectionMatrix), normalMatrix=Mat(normalMatrix))
        with np.errstate(all="ignore"):
            ns["main"]()
        return {k: np.asarray(ns[k], dtype=F) for k in ("vNormal", "vPosition", "vViewPosition", "gl_Position")}


class PlanetShader:
    """`planetFragmentShader` of the page, executable per fragment."""

    def __init__(self, html=jr.REFERENCE_HTML):
        with open(html, encoding="utf-8") as f:
            page = f.read()
        a = page.index("const planetFragmentShader = `") + len("const planetFragmentShader = `")
        self.glsl = page[a:page.index("`", a)]
        self.python_source = transliterate(self.glsl)
        self.ns = dict(_BUILTINS)
        with np.errstate(all="ignore"):
            exec(compile(self.python_source, "<planetFragmentShader idx:99-202 transliterated>", "exec"), self.ns)

    def fragment(self, vPosition, vNormal, vViewPosition, color1, color2, temperature, seed, sunPosition=(50.0, 50.0, 50.0)):
        """gl_FragColor (float32[4]) for one fragment: the varyings and uniforms of idx:99-110."""
        ns = self.ns
        ns.update(vPosition=vec(3, *vPosition), vNormal=vec(3, *vNormal), vViewPosition=vec(3, *vViewPosition), color1=vec(3, *color1),
                  color2=vec(3, *color2), temperature=F(temperature), seed=F(seed), sunPosition=vec(3, *sunPosition), time=F(0),
                  vUv=vec(2, 0.0, 0.0))
        with np.errstate(all="ignore"):
            ns["main"]()
        return np.asarray(ns["gl_FragColor"], dtype=F)

    def snoise(self, v):
        with np.errstate(all="ignore"):
            return F(self.ns["snoise"](vec(3, *v)))
