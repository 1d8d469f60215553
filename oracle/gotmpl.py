"""CPU restatement of the Go ``text/template`` + Sprig subset the reference chart uses.

TEST INFRASTRUCTURE ONLY (the checker): imported by ``tests/`` and by
``tests/golden/make_golden.py``; never by ``kvedge_b200``.

The reference's "program" is a Helm chart; what executes it is Helm v3 (Go
``text/template`` with ``github.com/Masterminds/sprig/v3`` functions), which is
third-party code that is NOT under /root/reference and cannot be run here (no
``helm``/``go`` in the image; version unpinned by the reference,
``deployment/helm/Chart.yaml:1`` only says ``apiVersion: v2``).  This module
restates the published semantics of exactly the constructs the chart relies on:

* actions / trim markers / comments ........ ``templates/_helper.tpl:1-75``
  (incl. the comment at ``_helper.tpl:21`` that itself contains ``{{ }}``)
* ``define`` / ``include`` ................. ``_helper.tpl:6,13,20,31``; all templates
* ``if`` / ``end`` ......................... ``_helper.tpl:22-24``; ``aziot-edge-vm-service.yaml:1,18``
* ``default`` ``trunc`` ``trimSuffix`` ..... ``_helper.tpl:7``
* ``printf`` ``replace`` ................... ``_helper.tpl:14`` (dead code)
* ``quote`` ................................ ``_helper.tpl:23``
* ``indent`` ............................... ``aziot-edge-data-volume-container.yaml:8``,
  ``aziot-edge-vm-service.yaml:6``, ``aziot-edge-vm.yaml:13``
* ``b64enc`` ............................... ``aziot-edge-runtime-config-secret.yaml:6``,
  ``aziot-edge-vm-cloud-init-secret.yaml:6``
* ``eq`` ................................... ``aziot-edge-vm-service.yaml:1``

A few neighbours (``else``, ``with``, ``range``, variables, ``nindent``,
``b64dec``, ``ne``/``not``/``and``/``or`` ...) are included so that the test
chart under ``tests/fixtures`` can exercise the engine more widely.

PARITY PIN: the only real-Helm output the reference records is the rendered
``NOTES.txt`` in ``deployment/az-iot-edge-k8s-kubevirt-ascii.cast:2294``; the
renderer is pinned to it (``tests/test_oracle_render.py``).  The two ``b64enc``
sites are "parity unpinned" by the reference itself.

Strings are handled as ``bytes`` end to end, as Go strings are byte strings:
``trunc`` slices bytes, ``--set-file`` content passes through verbatim.
"""
from __future__ import annotations

import re
from dataclasses import dataclass, field
from typing import Any, Callable, Dict, List, Optional, Tuple

from . import b64 as _b64


class TemplateError(Exception):
    pass


# --------------------------------------------------------------------------- values
class GoStruct:
    """A Go struct value: unknown field access is an error (unlike a map)."""

    def __init__(self, name: str, fields: Dict[str, Any]):
        self.name = name
        self.fields = fields


NOVALUE = object()  # what a missing map key evaluates to: prints as "<no value>"


def is_true(v: Any) -> bool:
    """text/template ``truth``: zero values are false."""
    if v is None or v is NOVALUE:
        return False
    if isinstance(v, bool):
        return v
    if isinstance(v, (int, float)):
        return v != 0
    if isinstance(v, (bytes, str, list, dict, tuple)):
        return len(v) > 0
    return True


def format_float(f: float) -> bytes:
    """Go ``%v`` of a float64 = strconv 'g' with the shortest round-trip digits; the
    exponent form is used when exp < -4 or exp >= 6 (shortest => eprec 6), which is
    why a values.yaml integer such as 1234567 (float64 after YAML->JSON) prints as
    1.234567e+06 in Helm."""
    if f != f:
        return b"NaN"
    if f in (float("inf"), float("-inf")):
        return b"+Inf" if f > 0 else b"-Inf"
    if f == 0:
        return b"-0" if str(f).startswith("-") else b"0"
    r = repr(float(f))
    mant, _, exp = r.partition("e")
    sign = ""
    if mant.startswith("-"):
        sign, mant = "-", mant[1:]
    ip, _, fp = mant.partition(".")
    if fp == "0":
        fp = ""
    e10 = int(exp) if exp else 0
    digits = (ip + fp).lstrip("0")
    # decimal point position relative to the digit string
    lead_zeros = len(ip + fp) - len((ip + fp).lstrip("0"))
    dp = len(ip) + e10 - lead_zeros
    digits = digits.rstrip("0") or "0"
    x = dp - 1
    if x < -4 or x >= 6:
        m = digits[0] + ("." + digits[1:] if len(digits) > 1 else "")
        return f"{sign}{m}e{'+' if x >= 0 else '-'}{abs(x):02d}".encode()
    if dp <= 0:
        return f"{sign}0.{'0' * (-dp)}{digits}".encode()
    if dp >= len(digits):
        return f"{sign}{digits}{'0' * (dp - len(digits))}".encode()
    return f"{sign}{digits[:dp]}.{digits[dp:]}".encode()


def go_print(v: Any) -> bytes:
    """``fmt.Fprint`` of one template value."""
    if v is NOVALUE:
        return b"<no value>"
    if v is None:
        return b"<nil>"
    if isinstance(v, bytes):
        return v
    if isinstance(v, str):
        return v.encode("utf-8")
    if isinstance(v, bool):
        return b"true" if v else b"false"
    if isinstance(v, int):
        return str(v).encode()
    if isinstance(v, float):
        return format_float(v)
    if isinstance(v, (list, tuple)):
        return b"[" + b" ".join(go_print(x) for x in v) + b"]"
    if isinstance(v, dict):
        items = sorted(v.items(), key=lambda kv: go_print(kv[0]))
        return b"map[" + b" ".join(go_print(k) + b":" + go_print(x) for k, x in items) + b"]"
    raise TemplateError(f"cannot print value of type {type(v).__name__}")


def strval(v: Any) -> bytes:
    """Sprig ``strval``."""
    if isinstance(v, bytes):
        return v
    if isinstance(v, str):
        return v.encode("utf-8")
    return go_print(v)


def string_arg(v: Any) -> bytes:
    """An argument bound to a Go ``string`` parameter.  text/template's validateType refuses
    anything that is not a string (``wrong type for value; expected string; got int64``;
    for nil or a missing key ``invalid value; expected string``): Sprig's b64enc, trunc,
    indent, trimSuffix, replace ... are plain ``func(string) string``.  Reference sites:
    aziot-edge-runtime-config-secret.yaml:6 (``--set azIotEdgeConfig=12345`` aborts in Helm)."""
    if isinstance(v, bytes):
        return v
    if isinstance(v, str):
        return v.encode("utf-8")
    if v is None or v is NOVALUE:
        raise TemplateError("invalid value; expected string")
    raise TemplateError("wrong type for value; expected string; got " + _go_typename(v).decode())


def go_quote(s: bytes) -> bytes:
    """``strconv.Quote``: double-quoted Go string literal of ``s``."""
    out = bytearray(b'"')
    i = 0
    n = len(s)
    simple = {0x07: b"\\a", 0x08: b"\\b", 0x0C: b"\\f", 0x0A: b"\\n", 0x0D: b"\\r",
              0x09: b"\\t", 0x0B: b"\\v", 0x5C: b"\\\\", 0x22: b'\\"'}
    while i < n:
        c = s[i]
        if c < 0x80:
            if c in simple:
                out += simple[c]
            elif c < 0x20 or c == 0x7F:
                out += b"\\x%02x" % c
            else:
                out.append(c)
            i += 1
            continue
        # multi-byte: decode one rune
        width = 0
        if 0xC2 <= c <= 0xDF:
            width = 2
        elif 0xE0 <= c <= 0xEF:
            width = 3
        elif 0xF0 <= c <= 0xF4:
            width = 4
        chunk = s[i:i + width] if width else b""
        try:
            ch = chunk.decode("utf-8") if width and len(chunk) == width else None
        except UnicodeDecodeError:
            ch = None
        if ch is None:
            out += b"\\x%02x" % c
            i += 1
            continue
        cp = ord(ch)
        if ch.isprintable() and cp != 0xFFFD or cp == 0xFFFD:
            out += chunk
        elif cp < 0x10000:
            out += b"\\u%04x" % cp
        else:
            out += b"\\U%08x" % cp
        i += width
    out += b'"'
    return bytes(out)


# --------------------------------------------------------------------------- lexer
_WS = b" \t\r\n"


@dataclass
class _Text:
    text: bytes


@dataclass
class _Action:
    src: str
    pos: int


def _lex(src: bytes, name: str) -> List[Any]:
    """Split template source into text and action items, applying ``{{-``/``-}}``
    trimming the way text/template's lexer does (a trim marker must be separated
    from the action body by white space)."""
    items: List[Any] = []
    i = 0
    n = len(src)
    trim_next = False
    while i < n:
        j = src.find(b"{{", i)
        if j < 0:
            j = n
        text = src[i:j]
        if trim_next:
            text = text.lstrip(_WS)
            trim_next = False
        if j >= n:
            if text:
                items.append(_Text(text))
            break
        k = j + 2
        ltrim = k + 1 < n and src[k:k + 1] == b"-" and src[k + 1:k + 2] in (b" ", b"\t", b"\r", b"\n")
        if ltrim:
            text = text.rstrip(_WS)
            k += 1
            while k < n and src[k:k + 1] in (b" ", b"\t", b"\r", b"\n"):
                k += 1
        if text:
            items.append(_Text(text))
        if src[k:k + 2] == b"/*":
            e = src.find(b"*/", k + 2)
            if e < 0:
                raise TemplateError(f"{name}: unclosed comment")
            e += 2
            if src[e:e + 2] == b"}}":
                i = e + 2
            elif src[e:e + 1] in (b" ", b"\t", b"\r", b"\n") and src[e + 1:e + 4] == b"-}}":
                i = e + 4
                trim_next = True
            else:
                raise TemplateError(f"{name}: comment ends before closing delimiter")
            continue
        # find the closing delimiter, skipping quoted strings
        p = k
        while True:
            if p >= n:
                raise TemplateError(f"{name}: unclosed action")
            c = src[p:p + 1]
            if c == b'"':
                p += 1
                while p < n and src[p:p + 1] != b'"':
                    if src[p:p + 1] == b"\\":
                        p += 1
                    if src[p:p + 1] == b"\n":
                        raise TemplateError(f"{name}: unterminated quoted string")
                    p += 1
                p += 1
                continue
            if c == b"`":
                q = src.find(b"`", p + 1)
                if q < 0:
                    raise TemplateError(f"{name}: unterminated raw quoted string")
                p = q + 1
                continue
            if c == b"'":
                p += 1
                while p < n and src[p:p + 1] != b"'":
                    if src[p:p + 1] == b"\\":
                        p += 1
                    p += 1
                p += 1
                continue
            if src[p:p + 2] == b"}}":
                break
            p += 1
        body = src[k:p]
        rtrim = len(body) >= 2 and body[-1:] == b"-" and body[-2:-1] in (b" ", b"\t", b"\r", b"\n")
        if rtrim:
            body = body[:-1].rstrip(_WS)
            trim_next = True
        items.append(_Action(body.decode("utf-8").strip(), j))
        i = p + 2
    return items


_TOKEN = re.compile(r"""
    (?P<ws>\s+)
  | (?P<str>"(?:[^"\\\n]|\\.)*")
  | (?P<raw>`[^`]*`)
  | (?P<char>'(?:[^'\\\n]|\\.)+')
  | (?P<decl>:=)
  | (?P<assign>=)
  | (?P<pipe>\|)
  | (?P<lp>\()
  | (?P<rp>\))
  | (?P<comma>,)
  | (?P<num>[+-]?(?:0[xX][0-9a-fA-F_]+|\d[\d_]*(?:\.\d*)?(?:[eE][+-]?\d+)?|\.\d+(?:[eE][+-]?\d+)?))
  | (?P<var>\$[A-Za-z0-9_]*)
  | (?P<field>(?:\.[A-Za-z_][A-Za-z0-9_]*)+)
  | (?P<dot>\.)
  | (?P<ident>[A-Za-z_][A-Za-z0-9_]*)
""", re.X)


def _unquote(tok: str) -> bytes:
    """Go interpreted string literal -> bytes."""
    body = tok[1:-1]
    out = bytearray()
    i = 0
    esc = {"a": 7, "b": 8, "f": 12, "n": 10, "r": 13, "t": 9, "v": 11, "\\": 92, '"': 34, "'": 39}
    while i < len(body):
        c = body[i]
        if c != "\\":
            out += c.encode("utf-8")
            i += 1
            continue
        i += 1
        c = body[i]
        if c in esc:
            out.append(esc[c])
            i += 1
        elif c == "x":
            out.append(int(body[i + 1:i + 3], 16))
            i += 3
        elif c == "u":
            out += chr(int(body[i + 1:i + 5], 16)).encode("utf-8")
            i += 5
        elif c == "U":
            out += chr(int(body[i + 1:i + 9], 16)).encode("utf-8")
            i += 9
        elif c in "01234567":
            out.append(int(body[i:i + 3], 8))
            i += 3
        else:
            raise TemplateError(f"unknown escape \\{c}")
    return bytes(out)


def _tokenize(s: str, where: str) -> List[Tuple[str, str]]:
    toks: List[Tuple[str, str]] = []
    i = 0
    while i < len(s):
        m = _TOKEN.match(s, i)
        if not m:
            raise TemplateError(f"{where}: unexpected {s[i:i + 10]!r} in action {{{{{s}}}}}")
        kind = m.lastgroup
        if kind == "ws":
            toks.append(("ws", " "))
        else:
            toks.append((kind, m.group(kind)))
        i = m.end()
    return toks


# --------------------------------------------------------------------------- AST
@dataclass
class Node:
    pass


@dataclass
class TextNode(Node):
    text: bytes


@dataclass
class Operand(Node):
    kind: str            # dot field var str num bool nil ident paren
    value: Any = None
    fields: List[str] = field(default_factory=list)


@dataclass
class Command(Node):
    args: List[Operand]


@dataclass
class Pipeline(Node):
    cmds: List[Command]
    decl: Optional[str] = None
    is_assign: bool = False


@dataclass
class ActionNode(Node):
    pipe: Pipeline


@dataclass
class IfNode(Node):
    kind: str  # if | with
    pipe: Pipeline
    body: List[Node]
    else_body: Optional[List[Node]]


@dataclass
class RangeNode(Node):
    pipe: Pipeline
    vars: List[str]
    body: List[Node]
    else_body: Optional[List[Node]]


@dataclass
class TemplateNode(Node):
    name: str
    pipe: Optional[Pipeline]


class _PipeParser:
    def __init__(self, toks: List[Tuple[str, str]], where: str):
        self.toks = toks
        self.i = 0
        self.where = where

    def peek(self, skip_ws=True):
        j = self.i
        while skip_ws and j < len(self.toks) and self.toks[j][0] == "ws":
            j += 1
        return self.toks[j] if j < len(self.toks) else ("eof", "")

    def next(self, skip_ws=True):
        while skip_ws and self.i < len(self.toks) and self.toks[self.i][0] == "ws":
            self.i += 1
        if self.i >= len(self.toks):
            return ("eof", "")
        t = self.toks[self.i]
        self.i += 1
        return t

    def pipeline(self, end: str = "eof") -> Pipeline:
        decl = None
        is_assign = False
        # variable declaration:  $x := pipeline   /  $x = pipeline
        save = self.i
        t = self.peek()
        if t[0] == "var":
            self.next()
            t2 = self.peek()
            if t2[0] in ("decl", "assign"):
                self.next()
                decl = t[1]
                is_assign = t2[0] == "assign"
            else:
                self.i = save
        cmds = [self.command(end)]
        while self.peek()[0] == "pipe":
            self.next()
            cmds.append(self.command(end))
        t = self.peek()
        if t[0] != end:
            raise TemplateError(f"{self.where}: unexpected {t[1]!r} in pipeline")
        return Pipeline(cmds, decl, is_assign)

    def command(self, end: str) -> Command:
        args: List[Operand] = []
        while True:
            t = self.peek()
            if t[0] in ("pipe", end, "eof", "rp"):
                break
            args.append(self.operand())
        if not args:
            raise TemplateError(f"{self.where}: missing value for command")
        return Command(args)

    def operand(self) -> Operand:
        kind, val = self.next()
        if kind == "dot":
            op = Operand("dot")
        elif kind == "field":
            op = Operand("field", fields=val[1:].split("."))
        elif kind == "var":
            op = Operand("var", val)
            nk = self.peek(skip_ws=False)
            if nk[0] == "field":
                self.next(skip_ws=False)
                op.fields = nk[1][1:].split(".")
        elif kind == "str":
            op = Operand("str", _unquote(val))
        elif kind == "raw":
            op = Operand("str", val[1:-1].encode("utf-8"))
        elif kind == "char":
            op = Operand("num", ord(_unquote(val).decode("utf-8")))
        elif kind == "num":
            txt = val.replace("_", "")
            try:
                op = Operand("num", int(txt, 0))
            except ValueError:
                op = Operand("num", float(txt))
        elif kind == "ident":
            if val == "true":
                op = Operand("bool", True)
            elif val == "false":
                op = Operand("bool", False)
            elif val == "nil":
                op = Operand("nil")
            else:
                op = Operand("ident", val)
        elif kind == "lp":
            inner = self.pipeline(end="rp")
            self.next()  # rp
            op = Operand("paren", inner)
            nk = self.peek(skip_ws=False)
            if nk[0] == "field":
                self.next(skip_ws=False)
                op.fields = nk[1][1:].split(".")
        else:
            raise TemplateError(f"{self.where}: unexpected {val!r} in operand")
        return op


_KEYWORD = re.compile(r"^(if|else|end|range|with|define|template|block|break|continue)\b\s*(.*)$", re.S)


class TemplateSet:
    """A namespace of named templates (Helm parses every file of a chart into one)."""

    def __init__(self, funcs: Optional[Dict[str, Callable]] = None):
        self.templates: Dict[str, List[Node]] = {}
        self.funcs: Dict[str, Callable] = dict(BUILTINS)
        self.funcs["include"] = self._include
        if funcs:
            self.funcs.update(funcs)

    # ---- parse
    def parse(self, name: str, src: bytes) -> None:
        items = _lex(src, name)
        pos = [0]

        def parse_pipe(text: str) -> Pipeline:
            return _PipeParser(_tokenize(text, name), name).pipeline()

        def parse_list(terminators: Tuple[str, ...]) -> Tuple[List[Node], str, str]:
            nodes: List[Node] = []
            while pos[0] < len(items):
                it = items[pos[0]]
                pos[0] += 1
                if isinstance(it, _Text):
                    nodes.append(TextNode(it.text))
                    continue
                m = _KEYWORD.match(it.src)
                if not m:
                    if it.src == "":
                        raise TemplateError(f"{name}: missing value for command")
                    nodes.append(ActionNode(parse_pipe(it.src)))
                    continue
                kw, rest = m.group(1), m.group(2).strip()
                if kw in ("end", "else"):
                    if kw not in terminators:
                        raise TemplateError(f"{name}: unexpected {{{{{kw}}}}}")
                    return nodes, kw, rest
                if kw in ("if", "with"):
                    nodes.append(parse_if(kw, rest))
                elif kw == "range":
                    nodes.append(parse_range(rest))
                elif kw == "define":
                    tname = _unquote(rest.strip()).decode("utf-8")
                    body, term, _ = parse_list(("end",))
                    if term != "end":
                        raise TemplateError(f"{name}: unexpected EOF in define")
                    self.templates[tname] = body
                elif kw == "template":
                    toks = _tokenize(rest, name)
                    pp = _PipeParser(toks, name)
                    k, v = pp.next()
                    if k not in ("str", "raw"):
                        raise TemplateError(f"{name}: template name must be a string")
                    tname = (_unquote(v) if k == "str" else v[1:-1].encode()).decode("utf-8")
                    pipe = None if pp.peek()[0] == "eof" else pp.pipeline()
                    nodes.append(TemplateNode(tname, pipe))
                else:
                    raise TemplateError(f"{name}: {{{{{kw}}}}} is not supported by this restatement")
            if terminators:
                raise TemplateError(f"{name}: unexpected EOF")
            return nodes, "", ""

        def parse_if(kw: str, rest: str) -> IfNode:
            pipe = parse_pipe(rest)
            body, term, trest = parse_list(("end", "else"))
            else_body = None
            if term == "else":
                m2 = re.match(r"^(if|with)\b\s*(.*)$", trest, re.S)
                if m2:
                    # {{ else if x }} == {{ else }}{{ if x }} ... {{ end }} sharing one end
                    else_body = [parse_if(m2.group(1), m2.group(2))]
                else:
                    else_body, term2, _ = parse_list(("end",))
            return IfNode(kw, pipe, body, else_body)

        def parse_range(rest: str) -> RangeNode:
            vars_: List[str] = []
            m2 = re.match(r"^(\$[A-Za-z0-9_]*)\s*(?:,\s*(\$[A-Za-z0-9_]*)\s*)?:=\s*(.*)$", rest, re.S)
            if m2:
                vars_ = [v for v in (m2.group(1), m2.group(2)) if v]
                rest = m2.group(3)
            pipe = parse_pipe(rest)
            body, term, _ = parse_list(("end", "else"))
            else_body = None
            if term == "else":
                else_body, _, _ = parse_list(("end",))
            return RangeNode(pipe, vars_, body, else_body)

        nodes, _, _ = parse_list(())
        self.templates[name] = nodes

    # ---- execute
    def execute(self, name: str, data: Any) -> bytes:
        if name not in self.templates:
            raise TemplateError(f"template {name!r} not defined")
        out = bytearray()
        self._root = data
        self._exec_list(self.templates[name], data, [{"$": data}], out)
        return bytes(out)

    def _include(self, name: Any, data: Any) -> bytes:
        """Helm's ``include``: execute a named template and return its output."""
        tname = strval(name).decode("utf-8")
        if tname not in self.templates:
            raise TemplateError(f"template {tname!r} not defined")
        out = bytearray()
        self._exec_list(self.templates[tname], data, [{"$": data}], out)
        return bytes(out)

    def _exec_list(self, nodes: List[Node], dot: Any, scopes: List[Dict[str, Any]], out: bytearray):
        for nd in nodes:
            if isinstance(nd, TextNode):
                out += nd.text
            elif isinstance(nd, ActionNode):
                v = self._eval_pipe(nd.pipe, dot, scopes)
                if nd.pipe.decl is None:
                    out += go_print(v)
            elif isinstance(nd, IfNode):
                v = self._eval_pipe(nd.pipe, dot, scopes + [{}])
                if is_true(v):
                    self._exec_list(nd.body, v if nd.kind == "with" else dot, scopes + [{}], out)
                elif nd.else_body is not None:
                    self._exec_list(nd.else_body, dot, scopes + [{}], out)
            elif isinstance(nd, RangeNode):
                v = self._eval_pipe(nd.pipe, dot, scopes + [{}])
                if isinstance(v, dict):
                    seq = [(k, v[k]) for k in sorted(v, key=lambda k: go_print(k))]
                elif isinstance(v, (list, tuple)):
                    seq = list(enumerate(v))
                elif v is None or v is NOVALUE:
                    seq = []
                elif isinstance(v, int) and not isinstance(v, bool):
                    seq = [(i, i) for i in range(v)]
                else:
                    raise TemplateError(f"range can't iterate over {go_print(v)!r}")
                if not seq and nd.else_body is not None:
                    self._exec_list(nd.else_body, dot, scopes + [{}], out)
                for k, x in seq:
                    sc: Dict[str, Any] = {}
                    if len(nd.vars) == 1:
                        sc[nd.vars[0]] = x
                    elif len(nd.vars) == 2:
                        sc[nd.vars[0]] = k
                        sc[nd.vars[1]] = x
                    self._exec_list(nd.body, x, scopes + [sc], out)
            elif isinstance(nd, TemplateNode):
                d = self._eval_pipe(nd.pipe, dot, scopes) if nd.pipe is not None else None
                if nd.name not in self.templates:
                    raise TemplateError(f"template {nd.name!r} not defined")
                self._exec_list(self.templates[nd.name], d, [{"$": d}], out)
            else:  # pragma: no cover
                raise TemplateError("unknown node")

    def _lookup_var(self, name: str, scopes: List[Dict[str, Any]]) -> Any:
        for sc in reversed(scopes):
            if name in sc:
                return sc[name]
        raise TemplateError(f"undefined variable {name!r}")

    def _field_chain(self, v: Any, fields: List[str]) -> Any:
        for f in fields:
            if isinstance(v, GoStruct):
                if f not in v.fields:
                    raise TemplateError(f"can't evaluate field {f} in type {v.name}")
                v = v.fields[f]
            elif isinstance(v, dict):
                v = v.get(f.encode("utf-8"), v.get(f, NOVALUE))
            elif v is None or v is NOVALUE:
                raise TemplateError(f"nil pointer evaluating interface {{}}.{f}")
            else:
                raise TemplateError(f"can't evaluate field {f} in type {type(v).__name__}")
        return v

    def _eval_operand(self, op: Operand, dot: Any, scopes, final: Any = NOVALUE, has_final=False) -> Any:
        if op.kind == "dot":
            return dot
        if op.kind == "field":
            return self._field_chain(dot, op.fields)
        if op.kind == "var":
            return self._field_chain(self._lookup_var(op.value, scopes), op.fields)
        if op.kind in ("str", "num", "bool"):
            return op.value
        if op.kind == "nil":
            return None
        if op.kind == "paren":
            return self._field_chain(self._eval_pipe(op.value, dot, scopes), op.fields)
        if op.kind == "ident":
            return self._call(op.value, [], dot, scopes)
        raise TemplateError("bad operand")

    def _call(self, fname: str, args: List[Any], dot, scopes) -> Any:
        if fname not in self.funcs:
            raise TemplateError(f"function {fname!r} not defined")
        return self.funcs[fname](*args)

    def _eval_cmd(self, cmd: Command, dot, scopes, piped: Any, has_piped: bool) -> Any:
        first = cmd.args[0]
        if first.kind == "ident":
            fname = first.value
            if fname in ("and", "or"):
                # short-circuit like Go >= 1.18 does; result is the deciding argument
                vals = cmd.args[1:]
                res: Any = None
                n_args = len(vals) + (1 if has_piped else 0)
                if n_args == 0:
                    raise TemplateError(f"wrong number of args for {fname}")
                for idx in range(n_args):
                    res = self._eval_operand(vals[idx], dot, scopes) if idx < len(vals) else piped
                    if (fname == "and") != is_true(res):
                        return res
                return res
            args = [self._eval_operand(a, dot, scopes) for a in cmd.args[1:]]
            if has_piped:
                args.append(piped)
            return self._call(fname, args, dot, scopes)
        if len(cmd.args) > 1 or has_piped:
            raise TemplateError("can't give argument to non-function")
        return self._eval_operand(first, dot, scopes)

    def _eval_pipe(self, pipe: Pipeline, dot, scopes) -> Any:
        v: Any = None
        has = False
        for cmd in pipe.cmds:
            v = self._eval_cmd(cmd, dot, scopes, v, has)
            has = True
        if pipe.decl is not None:
            if pipe.is_assign:
                for sc in reversed(scopes):
                    if pipe.decl in sc:
                        sc[pipe.decl] = v
                        break
                else:
                    raise TemplateError(f"undefined variable {pipe.decl!r}")
            else:
                scopes[-1][pipe.decl] = v
        return v


# --------------------------------------------------------------------------- functions
def _is_empty(v: Any) -> bool:
    """Sprig ``empty``: the zero value of the type."""
    return not is_true(v)


def _default(d: Any, *given: Any) -> Any:
    """Sprig ``default`` (reference: _helper.tpl:7)."""
    if not given or _is_empty(given[0]):
        return d
    return given[0]


def _toint(v: Any) -> int:
    if isinstance(v, bool):
        return int(v)
    if isinstance(v, (int, float)):
        return int(v)
    try:
        return int(strval(v))
    except ValueError:
        return 0


def _trunc(c: Any, s: Any) -> bytes:
    """Sprig ``trunc`` (reference: _helper.tpl:7,14). Slices BYTES, as Go does."""
    c = _toint(c)
    s = string_arg(s)
    if c < 0 and len(s) + c > 0:
        return s[len(s) + c:]
    if c >= 0 and len(s) > c:
        return s[:c]
    return s


def _trim_suffix(suffix: Any, s: Any) -> bytes:
    """Sprig ``trimSuffix`` = strings.TrimSuffix(s, suffix) (reference: _helper.tpl:7,14)."""
    suffix, s = string_arg(suffix), string_arg(s)
    if suffix and s.endswith(suffix):
        return s[:len(s) - len(suffix)]
    return s


def _trim_prefix(prefix: Any, s: Any) -> bytes:
    prefix, s = string_arg(prefix), string_arg(s)
    if prefix and s.startswith(prefix):
        return s[len(prefix):]
    return s


def _replace(old: Any, new: Any, src: Any) -> bytes:
    """Sprig ``replace`` = strings.Replace(src, old, new, -1) (reference: _helper.tpl:14)."""
    old, new, src = string_arg(old), string_arg(new), string_arg(src)
    if old == b"":
        # Go inserts `new` before every rune and at the end
        runes = src.decode("utf-8", errors="surrogateescape")
        parts = [c.encode("utf-8", errors="surrogateescape") for c in runes]
        return new + new.join(parts) + (new if parts else b"")
    return src.replace(old, new)


def _quote(*args: Any) -> bytes:
    """Sprig ``quote`` (reference: _helper.tpl:23): %q of every non-nil argument."""
    return b" ".join(go_quote(strval(a)) for a in args if a is not None and a is not NOVALUE)


def _squote(*args: Any) -> bytes:
    return b" ".join(b"'" + go_print(a) + b"'" for a in args if a is not None and a is not NOVALUE)


def _indent(n: Any, s: Any) -> bytes:
    """Sprig ``indent`` (reference: aziot-edge-vm.yaml:13 and three more sites):
    pad + strings.Replace(s, "\\n", "\\n"+pad, -1)."""
    pad = b" " * _toint(n)
    return pad + string_arg(s).replace(b"\n", b"\n" + pad)


def _nindent(n: Any, s: Any) -> bytes:
    return b"\n" + _indent(n, s)


def _b64enc(s: Any) -> bytes:
    """Sprig ``b64enc`` (reference: aziot-edge-runtime-config-secret.yaml:6,
    aziot-edge-vm-cloud-init-secret.yaml:6)."""
    return _b64.encode(string_arg(s))


def _b64dec(s: Any) -> bytes:
    """Sprig ``b64dec``: on error returns the error text, as Sprig does."""
    try:
        return _b64.decode(string_arg(s))
    except _b64.CorruptInputError as e:
        return str(e).encode()


def _basic_kind(v: Any) -> str:
    if isinstance(v, bool):
        return "bool"
    if isinstance(v, int):
        return "int"
    if isinstance(v, float):
        return "float"
    if isinstance(v, (bytes, str)):
        return "string"
    if v is None or v is NOVALUE:
        return "nil"
    return "other"


def _eq(*args: Any) -> bool:
    """text/template ``eq`` (reference: aziot-edge-vm-service.yaml:1): arg1 == any of the rest;
    comparing different basic kinds is an error."""
    if len(args) < 2:
        raise TemplateError("missing argument for comparison")
    a = args[0]
    ka = _basic_kind(a)
    for b in args[1:]:
        kb = _basic_kind(b)
        if ka == "nil" or kb == "nil":
            if ka == kb:
                return True
            continue
        if ka != kb:
            raise TemplateError("incompatible types for comparison")
        if ka == "other":
            raise TemplateError("invalid type for comparison")
        if (strval(a) == strval(b)) if ka == "string" else (a == b):
            return True
    return False


def _cmp(op: str):
    def f(a: Any, b: Any) -> bool:
        ka, kb = _basic_kind(a), _basic_kind(b)
        if ka != kb:
            raise TemplateError("incompatible types for comparison")
        if ka not in ("int", "float", "string"):
            raise TemplateError("invalid type for comparison")
        if ka == "string":
            a, b = strval(a), strval(b)
        return {"lt": a < b, "le": a <= b, "gt": a > b, "ge": a >= b}[op]
    return f


def _sprintf(fmt: Any, *args: Any) -> bytes:
    """Go ``fmt.Sprintf`` for the verbs charts use: %s %v %d %q %t %%."""
    fmt = strval(fmt)
    out = bytearray()
    i = 0
    ai = 0
    while i < len(fmt):
        c = fmt[i:i + 1]
        if c != b"%":
            out += c
            i += 1
            continue
        i += 1
        if i >= len(fmt):
            out += b"%!(NOVERB)"
            break
        v = fmt[i:i + 1]
        i += 1
        if v == b"%":
            out += b"%"
            continue
        if ai >= len(args):
            out += b"%!" + v + b"(MISSING)"
            continue
        a = args[ai]
        ai += 1
        if v in (b"s", b"v"):
            out += go_print(a)
        elif v == b"d":
            if isinstance(a, int) and not isinstance(a, bool):
                out += str(a).encode()
            else:
                out += b"%!d(" + _go_typename(a) + b"=" + go_print(a) + b")"
        elif v == b"q":
            out += go_quote(strval(a))
        elif v == b"t":
            out += go_print(a) if isinstance(a, bool) else b"%!t(" + _go_typename(a) + b"=" + go_print(a) + b")"
        else:
            raise TemplateError(f"printf verb %{v.decode()} is not supported by this restatement")
    if ai < len(args):
        out += b"%!(EXTRA " + b", ".join(_go_typename(a) + b"=" + go_print(a) for a in args[ai:]) + b")"
    return bytes(out)


def _go_typename(a: Any) -> bytes:
    return {"bool": b"bool", "int": b"int", "float": b"float64", "string": b"string",
            "nil": b"<nil>"}.get(_basic_kind(a), b"interface {}")


def _required(msg: Any, v: Any) -> Any:
    if v is None or v is NOVALUE or (isinstance(v, (bytes, str)) and len(v) == 0):
        raise TemplateError(strval(msg).decode("utf-8", "replace"))
    return v


def _len(v: Any) -> int:
    if isinstance(v, (bytes, str, list, tuple, dict)):
        return len(v)
    raise TemplateError("len of type " + type(v).__name__)


def _print(*args: Any) -> bytes:
    # fmt.Sprint adds spaces between operands when neither is a string
    out = bytearray()
    prev_str = True
    for i, a in enumerate(args):
        is_s = isinstance(a, (bytes, str))
        if i > 0 and not is_s and not prev_str:
            out += b" "
        out += go_print(a)
        prev_str = is_s
    return bytes(out)


BUILTINS: Dict[str, Callable] = {
    "default": _default,
    "trunc": _trunc,
    "trimSuffix": _trim_suffix,
    "trimPrefix": _trim_prefix,
    "replace": _replace,
    "quote": _quote,
    "squote": _squote,
    "indent": _indent,
    "nindent": _nindent,
    "b64enc": _b64enc,
    "b64dec": _b64dec,
    "eq": _eq,
    "ne": lambda a, b: not _eq(a, b),
    "lt": _cmp("lt"), "le": _cmp("le"), "gt": _cmp("gt"), "ge": _cmp("ge"),
    "not": lambda v: not is_true(v),
    "printf": _sprintf,
    "print": _print,
    "println": lambda *a: b" ".join(go_print(x) for x in a) + b"\n",
    "len": _len,
    "empty": _is_empty,
    "required": _required,
    "lower": lambda s: string_arg(s).decode("utf-8", "surrogateescape").lower().encode("utf-8", "surrogateescape"),
    "upper": lambda s: string_arg(s).decode("utf-8", "surrogateescape").upper().encode("utf-8", "surrogateescape"),
    "trim": lambda s: string_arg(s).decode("utf-8", "surrogateescape").strip().encode("utf-8", "surrogateescape"),
    "toString": lambda v: strval(v),
    "hasKey": lambda d, k: (strval(k) in d) or (strval(k).decode() in d),
    "ternary": lambda t, f, c: t if is_true(c) else f,
    "cat": lambda *a: b" ".join(strval(x) for x in a if x is not None and x is not NOVALUE),
    "contains": lambda sub, s: string_arg(sub) in string_arg(s),
    "hasPrefix": lambda p, s: string_arg(s).startswith(string_arg(p)),
    "hasSuffix": lambda p, s: string_arg(s).endswith(string_arg(p)),
    "repeat": lambda n, s: string_arg(s) * _toint(n),
    "sha256sum": lambda s: __import__("hashlib").sha256(string_arg(s)).hexdigest().encode(),
}
