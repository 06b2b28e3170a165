"""A small evaluator for the subset of Go/Helm templates the reference charts use.

Supported (SURVEY.md §2B "Go-template features actually used"): `.Release.Name`,
`.Release.Service`, `.Release.Time`, `.Chart.*`, `.Values.a.b` lookups, the dot inside
`range`, pipelines `a | f`, string / int literals, and the functions `tpl`, `int`, `gt`,
`ge`, `lt`, `le`, `eq`, `ne`, `and`, `or`, `not`, `add`, `sub`, `mul`, `div`, `required`,
`default`, `quote`, `toString`, `date` (Go reference-time layouts) plus the control actions
`if / else if / else / end`, `range / end`, `with / end`, template variables (`$x := v`, `$x = v`,
`range $v := ...`, `range $k, $v := ...`, `$x.field`, `$` for the root) and `{{-` / `-}}` whitespace
trimming - enough to evaluate the reference's chart templates themselves, which `tests/test_chart_templates.py`
does to hold `jobspec/render.py` against them.

E.g. `"{{ mul .Values.resources.gpu_nodes .Values.resources.gpus_per_node }}"`
(charts/machine-learning/training/mpijob-horovod-tensorflow-gpu/values.yaml:64-65) and
`"$HOME/logs/{{ .Release.Name }}-{{ date \"2006-01-02-15-04-05\" .Release.Time }}"` (:61-62).

This is an interpreter written for this project, not a port of text/template: it
tokenises actions, builds a tiny tree and walks it.
"""
from __future__ import annotations

import datetime as _dt
import re
from dataclasses import dataclass, field
from typing import Any


class TemplateError(ValueError):
    pass


_ACTION = re.compile(r"\{\{(-?)(.*?)(-?)\}\}", re.S)


@dataclass
class _Node:
    kind: str                      # text | expr | if | range | with | assign
    text: str = ""
    expr: str = ""
    branches: list = field(default_factory=list)   # if: [(cond_expr|None, [nodes])]
    body: list = field(default_factory=list)       # range body
    else_body: list = field(default_factory=list)
    names: tuple = ()              # range: ($key, $value) or ($value,); assign: ($name,)
    declare: bool = False          # assign: `:=` (new variable in this scope) vs `=` (nearest enclosing one)


def _lex(src: str):
    """Yields ('text', s) and ('action', expr) with whitespace trimming applied."""
    out, pos = [], 0
    for m in _ACTION.finditer(src):
        text = src[pos:m.start()]
        if m.group(1) == "-":
            text = text.rstrip()
        out.append(["text", text])
        body = m.group(2)
        # `{{- x -}}` markers must be followed/preceded by whitespace to count (Go rule)
        out.append(["action", body.strip(), m.group(3) == "-"])
        pos = m.end()
    out.append(["text", src[pos:]])
    # apply right-trim markers to the following text
    for i, tok in enumerate(out):
        if tok[0] == "action" and tok[2] and i + 1 < len(out):
            out[i + 1][1] = out[i + 1][1].lstrip()
    return [(t[0], t[1]) for t in out]


_ASSIGN = re.compile(r"^(\$\w+)\s*(:=|=)\s*(.+)$", re.S)
_RANGE_VARS = re.compile(r"^(\$\w+)\s*(?:,\s*(\$\w+)\s*)?:=\s*(.+)$", re.S)


def _parse(tokens, i=0, stop=()):
    nodes = []
    while i < len(tokens):
        kind, val = tokens[i]
        if kind == "text":
            if val:
                nodes.append(_Node("text", text=val))
            i += 1
            continue
        word = val.split(None, 1)[0] if val else ""
        if word in stop:
            return nodes, i
        if word == "if":
            node = _Node("if")
            cond = val[2:].strip()
            i += 1
            while True:
                body, i = _parse(tokens, i, stop=("else", "end"))
                node.branches.append((cond, body))
                if i >= len(tokens):
                    raise TemplateError("unterminated {{ if }}")
                w = tokens[i][1]
                i += 1
                if w.split(None, 1)[0] == "end":
                    break
                rest = w[4:].strip()          # after "else"
                if rest.startswith("if"):
                    cond = rest[2:].strip()
                else:
                    cond = None
            nodes.append(node)
        elif word in ("range", "with"):
            rest = val[len(word):].strip()
            node = _Node(word, expr=rest)
            m = _RANGE_VARS.match(rest) if word == "range" else _ASSIGN.match(rest)
            if m and word == "range":
                node.names = tuple(n for n in (m.group(1), m.group(2)) if n)
                node.expr = m.group(3).strip()
            elif m:
                node.names, node.expr = (m.group(1),), m.group(3).strip()
            node.body, i = _parse(tokens, i + 1, stop=("else", "end"))
            if i < len(tokens) and tokens[i][1].split(None, 1)[0] == "else":
                node.else_body, i = _parse(tokens, i + 1, stop=("end",))
            if i >= len(tokens):
                raise TemplateError("unterminated {{ range }}")
            i += 1
            nodes.append(node)
        elif word in ("end", "else"):
            raise TemplateError(f"unexpected {{{{ {word} }}}}")
        elif val.startswith("/*"):
            i += 1
        elif _ASSIGN.match(val):
            m = _ASSIGN.match(val)
            nodes.append(_Node("assign", expr=m.group(3).strip(), names=(m.group(1),), declare=m.group(2) == ":="))
            i += 1
        else:
            nodes.append(_Node("expr", expr=val))
            i += 1
    return nodes, i


# ------------------------------------------------------------------ expressions
_TOKEN = re.compile(r'\s*("(?:\\.|[^"\\])*"|`[^`]*`|\(|\)|\||[^\s()|]+)')


def _split(expr: str) -> list[str]:
    toks, pos = [], 0
    while pos < len(expr):
        m = _TOKEN.match(expr, pos)
        if not m:
            break
        toks.append(m.group(1))
        pos = m.end()
    return toks


_GO_LAYOUT = [("2006", "%Y"), ("01", "%m"), ("02", "%d"), ("15", "%H"), ("04", "%M"), ("05", "%S"),
              ("Jan", "%b"), ("Mon", "%a"), ("06", "%y"), ("03", "%I"), ("PM", "%p")]


def go_date(layout: str, t: _dt.datetime) -> str:
    fmt = layout
    for go, py in _GO_LAYOUT:
        fmt = fmt.replace(go, py)
    return t.strftime(fmt)


def _truthy(v: Any) -> bool:
    if v is None or v is False:
        return False
    if isinstance(v, (int, float)) and not isinstance(v, bool):
        return v != 0
    if isinstance(v, (str, list, dict, tuple)):
        return len(v) > 0
    return True


def _to_int(v: Any) -> int:
    if v is None or v == "":
        return 0
    if isinstance(v, bool):
        return int(v)
    try:
        return int(float(v))
    except (TypeError, ValueError):
        raise TemplateError(f"cannot convert {v!r} to int")


class Renderer:
    def __init__(self, context: dict):
        """context: {"Values": {...}, "Release": {...}, "Chart": {...}}"""
        self.ctx = context
        self.scopes: list[dict] = [{}]      # template variables ($x := ...), innermost scope last

    # -- public ---------------------------------------------------------------
    def render(self, src: str, dot: Any = None) -> str:
        if "{{" not in src:
            return src
        nodes, _ = _parse(_lex(src))
        return self._exec(nodes, self.ctx if dot is None else dot)

    # -- tree walk ------------------------------------------------------------
    def _exec(self, nodes, dot) -> str:
        out = []
        for n in nodes:
            if n.kind == "text":
                out.append(n.text)
            elif n.kind == "expr":
                v = self._eval(n.expr, dot)
                out.append(self._fmt(v))
            elif n.kind == "assign":
                self._set_var(n.names[0], self._eval(n.expr, dot), n.declare)
            elif n.kind == "if":
                for cond, body in n.branches:
                    if cond is None or _truthy(self._eval(cond, dot)):
                        out.append(self._scoped(body, dot))
                        break
            elif n.kind == "with":
                v = self._eval(n.expr, dot)
                if _truthy(v):
                    out.append(self._scoped(n.body, v, {n.names[0]: v} if n.names else None))
                else:
                    out.append(self._scoped(n.else_body, dot))
            elif n.kind == "range":
                seq = self._eval(n.expr, dot)
                if isinstance(seq, dict):       # Go ranges over maps in sorted key order
                    pairs = [(k, seq[k]) for k in sorted(seq, key=str)]
                elif isinstance(seq, int) and not isinstance(seq, bool):
                    pairs = [(k, k) for k in range(seq)]
                else:
                    pairs = list(enumerate(seq or []))
                if not pairs:
                    out.append(self._scoped(n.else_body, dot))
                for k, item in pairs:
                    names = {}
                    if len(n.names) == 2:
                        names = {n.names[0]: k, n.names[1]: item}
                    elif len(n.names) == 1:
                        names = {n.names[0]: item}
                    # with variables declared the dot stays what it was only in Go >= 1.x for `range $v :=`? No:
                    # the dot is still set to the element, and the charts rely on `$` for the root inside such loops
                    out.append(self._scoped(n.body, item, names))
        return "".join(out)

    def _scoped(self, body, dot, names: dict | None = None) -> str:
        self.scopes.append(dict(names or {}))
        try:
            return self._exec(body, dot)
        finally:
            self.scopes.pop()

    def _set_var(self, name: str, value: Any, declare: bool) -> None:
        if not declare:
            for sc in reversed(self.scopes):
                if name in sc:
                    sc[name] = value
                    return
            raise TemplateError(f"assignment to undeclared variable {name}")
        self.scopes[-1][name] = value

    def _get_var(self, name: str) -> Any:
        for sc in reversed(self.scopes):
            if name in sc:
                return sc[name]
        raise TemplateError(f"undefined variable {name}")

    @staticmethod
    def _fmt(v: Any) -> str:
        if v is None:
            return ""
        if isinstance(v, bool):
            return "true" if v else "false"
        if isinstance(v, float) and v.is_integer():
            return str(int(v))
        return str(v)

    # -- expression evaluation --------------------------------------------------
    def _eval(self, expr: str, dot) -> Any:
        toks = _split(expr)
        val, pos = self._pipeline(toks, 0, dot)
        if pos != len(toks):
            raise TemplateError(f"trailing tokens in {{{{ {expr} }}}}")
        return val

    def _pipeline(self, toks, pos, dot):
        val, pos = self._command(toks, pos, dot, piped=None)
        while pos < len(toks) and toks[pos] == "|":
            val, pos = self._command(toks, pos + 1, dot, piped=(val,))
        return val, pos

    def _command(self, toks, pos, dot, piped):
        if pos >= len(toks):
            raise TemplateError("empty command")
        head = toks[pos]
        if head in _FUNCS:
            args, pos = self._args(toks, pos + 1, dot)
            if piped is not None:
                args = args + list(piped)
            return _FUNCS[head](self, dot, *args), pos
        val, pos = self._operand(toks, pos, dot)
        return val, pos

    def _args(self, toks, pos, dot):
        args = []
        while pos < len(toks) and toks[pos] not in ("|", ")"):
            v, pos = self._operand(toks, pos, dot)
            args.append(v)
        return args, pos

    def _operand(self, toks, pos, dot):
        t = toks[pos]
        if t == "(":
            v, pos = self._pipeline(toks, pos + 1, dot)
            if pos >= len(toks) or toks[pos] != ")":
                raise TemplateError("missing )")
            return v, pos + 1
        if t.startswith('"'):
            return bytes(t[1:-1], "utf-8").decode("unicode_escape"), pos + 1
        if t.startswith("`"):
            return t[1:-1], pos + 1
        if re.fullmatch(r"-?\d+", t):
            return int(t), pos + 1
        if re.fullmatch(r"-?\d+\.\d*", t):
            return float(t), pos + 1
        if t in ("true", "false"):
            return t == "true", pos + 1
        if t == "nil":
            return None, pos + 1
        if t == ".":
            return dot, pos + 1
        if t == "$":
            return self.ctx, pos + 1
        if re.fullmatch(r"\$\w+(\.\w+)*", t):           # $name or $name.field.sub
            parts = t.split(".")
            cur = self._get_var(parts[0])
            for part in parts[1:]:
                cur = cur.get(part) if isinstance(cur, dict) else getattr(cur, part, None)
                if cur is None:
                    break
            return cur, pos + 1
        if t.startswith(".") or t.startswith("$."):
            root = self.ctx if (t.startswith("$.") or isinstance(dot, (str, int, float)) or dot is None) else dot
            path = t[2:] if t.startswith("$.") else t[1:]
            cur = root
            parts = path.split(".")
            # `.Values...` / `.Release...` always resolve from the root context
            if parts[0] in ("Values", "Release", "Chart", "Capabilities"):
                cur = self.ctx
            for part in parts:
                if isinstance(cur, dict):
                    cur = cur.get(part)
                else:
                    cur = getattr(cur, part, None)
                if cur is None:
                    break
            return cur, pos + 1
        if t in _FUNCS:  # zero-arg function used as an operand
            return _FUNCS[t](self, dot), pos + 1
        raise TemplateError(f"cannot evaluate token {t!r}")


def _f_required(r, dot, msg, val=None):
    if val is None or val == "":
        raise TemplateError(str(msg))
    return val


def _f_date(r, dot, layout, t):
    if not isinstance(t, _dt.datetime):
        t = _dt.datetime.now()
    return go_date(layout, t)


def _f_default(r, dot, dflt, val=None):
    return val if _truthy(val) else dflt


_FUNCS = {
    "tpl": lambda r, dot, s, ctx=None: r.render(str(s)) if s is not None else "",
    "int": lambda r, dot, v: _to_int(v),
    "int64": lambda r, dot, v: _to_int(v),
    "float64": lambda r, dot, v: float(v),
    "toString": lambda r, dot, v: Renderer._fmt(v),
    "quote": lambda r, dot, *v: " ".join('"' + Renderer._fmt(x).replace('"', '\\"') + '"' for x in v),
    "squote": lambda r, dot, *v: " ".join("'" + Renderer._fmt(x) + "'" for x in v),
    "add": lambda r, dot, *v: sum(_to_int(x) for x in v),
    "sub": lambda r, dot, a, b: _to_int(a) - _to_int(b),
    "mul": lambda r, dot, *v: __import__("math").prod(_to_int(x) for x in v),
    "div": lambda r, dot, a, b: _to_int(a) // _to_int(b),
    "mod": lambda r, dot, a, b: _to_int(a) % _to_int(b),
    "gt": lambda r, dot, a, b: a > b,
    "ge": lambda r, dot, a, b: a >= b,
    "lt": lambda r, dot, a, b: a < b,
    "le": lambda r, dot, a, b: a <= b,
    "eq": lambda r, dot, a, *b: any(a == x for x in b),
    "ne": lambda r, dot, a, b: a != b,
    "and": lambda r, dot, *v: next((x for x in v if not _truthy(x)), v[-1]),
    "or": lambda r, dot, *v: next((x for x in v if _truthy(x)), v[-1]),
    "not": lambda r, dot, v: not _truthy(v),
    "required": _f_required,
    "default": _f_default,
    "date": _f_date,
    "now": lambda r, dot: _dt.datetime.now(),
    "len": lambda r, dot, v: len(v) if v is not None else 0,
    "lower": lambda r, dot, v: str(v).lower(),
    "upper": lambda r, dot, v: str(v).upper(),
    "trim": lambda r, dot, v: str(v).strip(),
    "toYaml": lambda r, dot, v: __import__("yaml").safe_dump(v, default_flow_style=False).rstrip(),
    "indent": lambda r, dot, n, s: "\n".join(" " * int(n) + l for l in str(s).split("\n")),
    "nindent": lambda r, dot, n, s: "\n" + "\n".join(" " * int(n) + l for l in str(s).split("\n")),
}


def render(src: str, values: dict, release_name: str, *, chart: dict | None = None,
           release_time: _dt.datetime | None = None) -> str:
    ctx = {"Values": values,
           "Release": {"Name": release_name, "Service": "b200flow", "Namespace": "local",
                       "Time": release_time or _dt.datetime.now()},
           "Chart": chart or {}}
    return Renderer(ctx).render(src)
