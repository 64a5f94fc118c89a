"""Cypher-subset front end (host side, cold path): query text -> plan.

The reference hands `run_query` Cypher TEXT (common/neo4j_query_executor.py:15-24).  This module
parses the subset its statements (and LLM-written variants of them) use and recognises the shapes
the engine executes:

  NodeScan  MATCH (n[:L]) WHERE <props> RETURN ...                    S1 find_srckind_metapath_neo4j.py:64-68,
                                                                      S6 analyze_root_cause.py:201-206
  VarLen    MATCH path = (a)-[*m..n]-(b) WHERE a.kind = .. AND ...    S3a/b :95-116; fixed-length untyped
            patterns with kind tests                                   S3c/d :119-129
  Chain     MATCH (a:K1)-[r:T]->(b:K2) WHERE r.key = .. (several)     S2 :76-84, S4 generate_query.py:223-260
  State     MATCH (n:K)-[r:HasState]->(s:KK) WHERE n.id = .. AND      S5/S5' analyze_root_cause.py:51-79
            r.tmin <= .. AND r.tmax > ..

A Chain statement may be written the way an assistant writes it rather than the way the generator does: one long
pattern or comma-separated ones, hops without a direction (the labels / kinds decide which way each relationship is
stored), unnamed nodes and relationships, `{prop: value}` maps, RETURN *, comments.  Its WHERE may hold any predicate:
`r.key = ..` picks the step's edge types, predicates on the seed node written before its LIMIT choose the seeds, a
kind test such as `pod.kind2 = 'Pod'` stands in for a label, and whatever is left (IN, STARTS WITH, =~, IS NULL,
type(), labels(), toLower() ...) is evaluated on the finished rows, host side.

Anything that does not parse, or parses but is not one of these shapes, raises CypherSyntaxError -
the exception the reference's retry loop catches (test_all.py:109).
"""
from __future__ import annotations

import re
import threading
from dataclasses import dataclass, field


class CypherSyntaxError(Exception):
    """Stands in for neo4j.exceptions.CypherSyntaxError (the shim module re-exports this class)."""

    def __init__(self, message, code="Neo.ClientError.Statement.SyntaxError"):
        super().__init__(message)
        self.message = message
        self.code = code


# ------------------------------------------------------------------------------------------ lexer
_KEYWORDS = {"MATCH", "WHERE", "WITH", "RETURN", "LIMIT", "DISTINCT", "AND", "OR", "NOT", "IN", "IS", "NULL",
             "CONTAINS", "AS", "STARTS", "ENDS", "TRUE", "FALSE", "OPTIONAL", "ORDER", "BY", "SKIP", "UNWIND",
             "CREATE", "MERGE", "DELETE", "SET", "CALL"}
_TOKEN = re.compile(r"""
    (?P<ws>\s+|//[^\n]*|/\*.*?\*/)
  | (?P<num>\d+\.\d+|\d+(?!\.\d))
  | (?P<str>'(?:[^'\\]|\\.)*'|"(?:[^"\\]|\\.)*")
  | (?P<param>\$[A-Za-z_][A-Za-z_0-9]*)
  | (?P<ident>[A-Za-z_][A-Za-z_0-9]*|`[^`]+`)
  | (?P<op><>|<=|>=|=~|->|<-|\.\.|[()\[\]{},.:;=<>\-*|+])
""", re.X | re.S)
_AGGREGATES = {"count", "collect", "sum", "avg", "min", "max"}
_ESC = {"n": "\n", "t": "\t", "r": "\r", "\\": "\\", "'": "'", '"': '"', "b": "\b", "f": "\f"}


def _unquote(s):
    if "\\" not in s:                                 # nothing to unescape: the usual case (ids, timestamps, keys, messages)
        return s[1:-1]
    out, i, body = [], 0, s[1:-1]
    while i < len(body):
        c = body[i]
        if c == "\\" and i + 1 < len(body):
            n = body[i + 1]
            if n == "u" and i + 5 < len(body) + 0:
                out.append(chr(int(body[i + 2:i + 6], 16)))
                i += 6
                continue
            if n == "x" and i + 3 < len(body) + 0:
                out.append(chr(int(body[i + 2:i + 4], 16)))
                i += 4
                continue
            out.append(_ESC.get(n, n))
            i += 2
        else:
            out.append(c)
            i += 1
    return "".join(out)


def tokenize(text):
    toks, pos = [], 0
    while pos < len(text):
        m = _TOKEN.match(text, pos)
        if not m:
            raise CypherSyntaxError(f"Invalid input {text[pos]!r} at offset {pos}: {text[max(0, pos - 20):pos + 20]!r}")
        pos = m.end()
        kind = m.lastgroup
        if kind == "ws":
            continue
        val = m.group()
        if kind == "ident":
            if val.startswith("`"):
                toks.append(("ident", val[1:-1], m.start()))
            elif val.upper() in _KEYWORDS:
                toks.append(("kw", val.upper(), m.start()))
            else:
                toks.append(("ident", val, m.start()))
        elif kind == "str":
            toks.append(("str", _unquote(val), m.start()))
        elif kind == "num":
            toks.append(("num", float(val) if "." in val else int(val), m.start()))
        elif kind == "param":
            toks.append(("param", val[1:], m.start()))
        else:
            toks.append(("op", val, m.start()))
    toks.append(("eof", None, len(text)))
    return toks


# ------------------------------------------------------------------------------------------ AST
@dataclass
class NodePat:
    var: str | None
    labels: list
    props: dict


@dataclass
class RelPat:
    var: str | None
    types: list
    direction: str            # 'out' (a)->(b), 'in' (a)<-(b), 'und'
    varlen: tuple | None      # (min, max) or None
    props: dict


@dataclass
class Pattern:
    path_var: str | None
    nodes: list
    rels: list


@dataclass
class Clause:
    kind: str                 # MATCH | WITH | RETURN
    patterns: list = field(default_factory=list)
    where: object = None
    items: list = field(default_factory=list)     # (expr, alias or None, source text)
    distinct: bool = False
    limit: int | None = None
    order: list = field(default_factory=list)      # ORDER BY: (expr, descending)
    skip: int = 0


class Parser:
    def __init__(self, text):
        self.text = text
        self.toks = tokenize(text)
        self.i = 0

    # -- helpers
    def peek(self, k=0):
        return self.toks[min(self.i + k, len(self.toks) - 1)]

    def err(self, what):
        kind, val, pos = self.peek()
        got = "end of input" if kind == "eof" else repr(val)
        raise CypherSyntaxError(f"Invalid input {got}: expected {what} (offset {pos}) "
                                f"\"{self.text[max(0, pos - 30):pos + 30].strip()}\"")

    def at(self, kind, val=None):
        t = self.peek()
        return t[0] == kind and (val is None or t[1] == val)

    def accept(self, kind, val=None):
        if self.at(kind, val):
            self.i += 1
            return self.toks[self.i - 1]
        return None

    def expect(self, kind, val=None, what=None):
        t = self.accept(kind, val)
        if t is None:
            self.err(what or (val if val is not None else kind))
        return t

    # -- query
    def parse(self):
        clauses = []
        while not self.at("eof"):
            if self.accept("op", ";"):
                continue
            if self.accept("kw", "MATCH"):
                c = Clause("MATCH")
                c.patterns.append(self.pattern())
                while self.accept("op", ","):
                    c.patterns.append(self.pattern())
                if self.accept("kw", "WHERE"):
                    c.where = self.expr()
                clauses.append(c)
            elif self.at("kw", "WITH") or self.at("kw", "RETURN"):
                c = Clause(self.peek()[1])
                self.i += 1
                c.distinct = bool(self.accept("kw", "DISTINCT"))
                c.items.append(self.ret_item())
                while self.accept("op", ","):
                    c.items.append(self.ret_item())
                if self.accept("kw", "ORDER"):
                    self.expect("kw", "BY", "BY after ORDER")
                    while True:
                        e = self.expr()
                        t = self.accept("ident")
                        if t is not None and t[1].upper() not in ("ASC", "ASCENDING", "DESC", "DESCENDING"):
                            self.i -= 1
                            self.err("ASC, DESC, SKIP, LIMIT or the end of the clause")
                        c.order.append((e, t is not None and t[1].upper().startswith("DESC")))
                        if not self.accept("op", ","):
                            break
                if self.accept("kw", "SKIP"):
                    c.skip = int(self.expect("num", what="an integer after SKIP")[1])
                if self.accept("kw", "LIMIT"):
                    c.limit = int(self.expect("num", what="an integer after LIMIT")[1])
                if c.kind == "WITH" and (c.order or c.skip):
                    raise _unsupported("ORDER BY / SKIP in a WITH clause")
                clauses.append(c)
            elif self.at("kw"):
                raise CypherSyntaxError(f"The clause {self.peek()[1]} is not supported by this executor "
                                        f"(read-only MATCH / WITH / RETURN statements only)")
            else:
                self.err("MATCH, WITH or RETURN")
        if not clauses:
            raise CypherSyntaxError("Unexpected end of input: expected MATCH")
        if clauses[-1].kind != "RETURN":
            raise CypherSyntaxError("Query cannot conclude with " + clauses[-1].kind +
                                    " (must be a RETURN clause)")
        return clauses

    def ret_item(self):
        if self.accept("op", "*"):                   # RETURN *: every named variable in scope (resolved per row)
            return (("star",), None, "*")
        start = self.peek()[2]
        e = self.expr()
        end = self.peek()[2]
        alias = None
        if self.accept("kw", "AS"):
            alias = self.expect("ident", what="an alias after AS")[1]
        return (e, alias, self.text[start:end].strip())

    # -- patterns
    def pattern(self):
        path_var = None
        if self.at("ident") and self.peek(1) == ("op", "=", self.peek(1)[2]):
            path_var = self.accept("ident")[1]
            self.expect("op", "=")
        nodes, rels = [self.node_pat()], []
        while self.at("op", "-") or self.at("op", "<-"):
            rels.append(self.rel_pat())
            nodes.append(self.node_pat())
        return Pattern(path_var, nodes, rels)

    def prop_map(self):
        props = {}
        if self.accept("op", "{"):
            while True:
                k = self.expect("ident", what="a property key")[1]
                self.expect("op", ":")
                props[k] = self.expr()
                if not self.accept("op", ","):
                    break
            self.expect("op", "}")
        return props

    def node_pat(self):
        self.expect("op", "(", "a node pattern '('")
        var = self.accept("ident")
        labels = []
        while self.accept("op", ":"):
            labels.append(self.expect("ident", what="a label")[1])
        props = self.prop_map()
        self.expect("op", ")", "')' closing the node pattern")
        return NodePat(var[1] if var else None, labels, props)

    def rel_pat(self):
        left_arrow = bool(self.accept("op", "<-"))
        if not left_arrow:
            self.expect("op", "-")
        var, types, varlen, props = None, [], None, {}
        if self.accept("op", "["):
            v = self.accept("ident")
            var = v[1] if v else None
            if self.accept("op", ":"):
                types.append(self.expect("ident", what="a relationship type")[1])
                while self.accept("op", "|"):
                    self.accept("op", ":")
                    types.append(self.expect("ident", what="a relationship type")[1])
            if self.accept("op", "*"):
                lo, hi = 1, None
                n = self.accept("num")
                if n:
                    lo = hi = int(n[1])
                if self.accept("op", ".."):
                    hi = None
                    m = self.accept("num")
                    if m:
                        hi = int(m[1])
                    if not n:
                        lo = 1
                varlen = (lo, hi)
            props = self.prop_map()
            self.expect("op", "]", "']' closing the relationship pattern")
        if self.accept("op", "->"):
            if left_arrow:
                raise CypherSyntaxError("A relationship pattern cannot point in both directions")
            direction = "out"
        else:
            self.expect("op", "-", "'-' or '->' after the relationship pattern")
            direction = "in" if left_arrow else "und"
        return RelPat(var, types, direction, varlen, props)

    # -- expressions: ('lit', v) ('param', n) ('var', n) ('prop', e, k) ('list', [..]) ('cmp', op, a, b)
    #    ('and', a, b) ('or', a, b) ('not', a) ('isnull', e, negated) ('call', name, [args])
    #    ('quant', kind, var, list_expr, pred) ('slice', e, lo, hi) ('index', e, i) ('neg', e)
    def expr(self):
        e = self.and_expr()
        while self.accept("kw", "OR"):
            e = ("or", e, self.and_expr())
        return e

    def and_expr(self):
        e = self.not_expr()
        while self.accept("kw", "AND"):
            e = ("and", e, self.not_expr())
        return e

    def not_expr(self):
        if self.accept("kw", "NOT"):
            return ("not", self.not_expr())
        return self.cmp_expr()

    def cmp_expr(self):
        a = self.unary()
        t = self.peek()
        if t[0] == "op" and t[1] in ("=", "<>", "<", "<=", ">", ">="):
            self.i += 1
            return ("cmp", t[1], a, self.unary())
        if t[0] == "op" and t[1] == "=~":
            self.i += 1
            return ("cmp", "regex", a, self.unary())
        if self.accept("kw", "IN"):
            return ("cmp", "in", a, self.unary())
        if self.accept("kw", "CONTAINS"):
            return ("cmp", "contains", a, self.unary())
        if self.at("kw", "STARTS") or self.at("kw", "ENDS"):
            op = "startswith" if self.peek()[1] == "STARTS" else "endswith"
            self.i += 1
            self.expect("kw", "WITH")
            return ("cmp", op, a, self.unary())
        if self.accept("kw", "IS"):
            neg = bool(self.accept("kw", "NOT"))
            self.expect("kw", "NULL")
            return ("isnull", a, neg)
        return a

    def unary(self):
        if self.accept("op", "-"):
            return ("neg", self.unary())
        return self.postfix()

    def postfix(self):
        e = self.primary()
        while True:
            if self.accept("op", "."):
                e = ("prop", e, self.expect("ident", what="a property name")[1])
            elif self.at("op", "["):
                self.i += 1
                lo = None if self.at("op", "..") else self.expr()
                if self.accept("op", ".."):
                    hi = None if self.at("op", "]") else self.expr()
                    e = ("slice", e, lo, hi)
                else:
                    e = ("index", e, lo)
                self.expect("op", "]")
            else:
                return e

    def primary(self):
        t = self.peek()
        if t[0] in ("str", "num"):
            self.i += 1
            return ("lit", t[1])
        if t[0] == "param":
            self.i += 1
            return ("param", t[1])
        if t[0] == "kw" and t[1] in ("TRUE", "FALSE", "NULL"):
            self.i += 1
            return ("lit", {"TRUE": True, "FALSE": False, "NULL": None}[t[1]])
        if self.accept("op", "["):
            items = []
            if not self.at("op", "]"):
                items.append(self.expr())
                while self.accept("op", ","):
                    items.append(self.expr())
            self.expect("op", "]")
            return ("list", items)
        if self.accept("op", "("):
            e = self.expr()
            self.expect("op", ")")
            return e
        if t[0] == "ident":
            self.i += 1
            if self.at("op", "("):
                name = t[1].lower()
                self.i += 1
                if name in ("all", "any", "single", "none"):
                    var = self.expect("ident", what="a variable")[1]
                    self.expect("kw", "IN")
                    lst = self.expr()
                    self.expect("kw", "WHERE")
                    pred = self.expr()
                    self.expect("op", ")")
                    return ("quant", name, var, lst, pred)
                args = []
                if name in _AGGREGATES:
                    raise _unsupported(f"aggregation ({t[1]}(...)): this executor returns the matched rows themselves")
                if not self.at("op", ")"):
                    args.append(self.expr())
                    while self.accept("op", ","):
                        args.append(self.expr())
                self.expect("op", ")")
                return ("call", name, args)
            return ("var", t[1])
        self.err("an expression")


def parse(text):
    if not isinstance(text, str) or not text.strip():
        raise CypherSyntaxError("Unexpected end of input: expected MATCH")
    return Parser(text).parse()


# The statements of one incident differ from those of the next only in their string literals (the message,
# ids, timestamps - the reference interpolates them into the text, analyze_root_cause.py:70-79).  parse_shared
# lifts the literals out as parameters and parses each distinct skeleton once.
_STR_LIT = re.compile(r"""'(?:[^'\\]|\\.)*'|"(?:[^"\\]|\\.)*["]""")
_RETURN_KW = re.compile(r"\bRETURN\b", re.I)
_SKELETONS = {}
# ... and compiled once per skeleton and per set of VALUES the compilation looked at (`_lit` notes every parameter it
# resolves: relationship keys, kinds - the same from incident to incident; the message, which stays an expression of
# the plan, is not among them).  A state statement's id and timestamps are resolved per call (StatePlan.late).
_PLANS = {}
_tls = threading.local()                             # .noted: the (name, value) pairs `_lit` resolved during one compilation


def parse_shared(text, with_skeleton=False):
    """-> (clauses, {parameter name: literal value}) [, skeleton text or None].  Falls back to a plain parse (and
    its error messages, which quote the text) whenever lifting is not clearly safe."""
    out = _parse_shared(text)
    return out if with_skeleton else out[:2]


def _parse_shared(text):
    if not isinstance(text, str) or "'" not in text and '"' not in text or "`" in text or "$__s" in text:
        return parse(text), {}, None
    lits = []

    def lift(m):
        lits.append(m.group())
        return f"$__s{len(lits) - 1}"

    skel = _STR_LIT.sub(lift, text)
    if "/*" in skel or "//" in skel:
        # a comment: a quote inside it is not a literal, and the lifting above cannot tell.  (A comment's opener always
        # survives in the skeleton - it stands before any quote of the comment - while `//` inside a real literal,
        # a URL in a message, has been lifted out and does not get here.)
        return parse(text), {}, None
    last_ret = None
    for last_ret in _RETURN_KW.finditer(skel):
        pass
    if last_ret is not None and "$__s" in skel[last_ret.end():]:
        return parse(text), {}, None                     # a literal in RETURN names its column: keep the text
    clauses = _SKELETONS.get(skel)
    if clauses is None:
        try:
            clauses = Parser(skel).parse()
        except CypherSyntaxError:
            return parse(text), {}, None                 # report the error against the statement as written
        if len(_SKELETONS) >= 512:
            _SKELETONS.clear()
            _PLANS.clear()
        _SKELETONS[skel] = clauses
    return clauses, {f"__s{i}": _unquote(v) for i, v in enumerate(lits)}, skel


# ------------------------------------------------------------------------------------------ evaluation
class _Null:
    pass


def eval_expr(e, env, params):
    """Three-valued evaluation of an expression over `env` (variable -> entity / value)."""
    k = e[0]
    if k == "lit":
        return e[1]
    if k == "param":
        if e[1] not in params:
            raise CypherSyntaxError(f"Expected parameter(s): {e[1]}", "Neo.ClientError.Statement.ParameterMissing")
        return params[e[1]]
    if k == "var":
        if e[1] not in env:
            raise CypherSyntaxError(f"Variable `{e[1]}` not defined")
        return env[e[1]]
    if k == "prop":
        base = eval_expr(e[1], env, params)
        if base is None:
            return None
        if isinstance(base, dict):
            return base.get(e[2])
        return base[e[2]]
    if k == "list":
        return [eval_expr(x, env, params) for x in e[1]]
    if k == "neg":
        v = eval_expr(e[1], env, params)
        return None if v is None else -v
    if k == "not":
        v = eval_expr(e[1], env, params)
        return None if v is None else (not v)
    if k == "and":
        a, b = eval_expr(e[1], env, params), eval_expr(e[2], env, params)
        if a is False or b is False:
            return False
        return None if a is None or b is None else True
    if k == "or":
        a, b = eval_expr(e[1], env, params), eval_expr(e[2], env, params)
        if a is True or b is True:
            return True
        return None if a is None or b is None else False
    if k == "isnull":
        v = eval_expr(e[1], env, params)
        return (v is not None) if e[2] else (v is None)
    if k == "cmp":
        a, b = eval_expr(e[2], env, params), eval_expr(e[3], env, params)
        op = e[1]
        if a is None or b is None:
            return None
        try:
            if op == "=":
                return a == b
            if op == "<>":
                return a != b
            if op == "in":
                return a in b
            if op == "contains":
                return b in a
            if op == "startswith":
                return a.startswith(b)
            if op == "endswith":
                return a.endswith(b)
            if op == "regex":                        # =~ matches the WHOLE string
                return re.fullmatch(b, a, re.S) is not None
            return {"<": a < b, "<=": a <= b, ">": a > b, ">=": a >= b}[op]
        except (TypeError, AttributeError, re.error):
            return None
    if k == "call":
        name, args = e[1], [eval_expr(x, env, params) for x in e[2]]
        fn = _FUNCTIONS.get(name)
        if fn is None:
            raise CypherSyntaxError(f"Unknown function '{name}'")
        if len(args) not in fn[0]:
            raise CypherSyntaxError(f"Insufficient parameters for function '{name}'" if len(args) < min(fn[0])
                                    else f"Too many parameters for function '{name}'")
        if name == "coalesce":
            return next((a for a in args if a is not None), None)
        if name == "exists":
            return args[0] is not None
        if args and args[0] is None:
            return None
        try:
            return fn[1](*args)
        except (TypeError, ValueError, AttributeError, IndexError):
            return None
    raise CypherSyntaxError(f"Expression of kind {k!r} is not supported here")


def _to_int(v):
    return int(float(v)) if isinstance(v, str) else int(v)


def _to_bool(v):
    if isinstance(v, str):
        return {"true": True, "false": False}.get(v.strip().lower())
    return bool(v)


# scalar functions an LLM-written statement is likely to wrap a property in: name -> (arities, implementation);
# a null first argument gives null (handled by the caller), an argument of the wrong type gives null as well
_FUNCTIONS = {
    "size": ((1,), len), "length": ((1,), len),
    "tolower": ((1,), lambda v: v.lower()), "toupper": ((1,), lambda v: v.upper()),
    "trim": ((1,), lambda v: v.strip()), "ltrim": ((1,), lambda v: v.lstrip()), "rtrim": ((1,), lambda v: v.rstrip()),
    "tostring": ((1,), lambda v: ("true" if v else "false") if isinstance(v, bool) else str(v)),
    "tointeger": ((1,), _to_int), "tofloat": ((1,), float), "toboolean": ((1,), _to_bool),
    "replace": ((3,), lambda v, a, b: v.replace(a, b)), "split": ((2,), lambda v, d: v.split(d)),
    "substring": ((2, 3), lambda v, a, n=None: v[a:] if n is None else v[a:a + n]),
    "left": ((2,), lambda v, n: v[:n]), "right": ((2,), lambda v, n: v[len(v) - n:] if n else ""),
    "head": ((1,), lambda v: v[0] if v else None), "last": ((1,), lambda v: v[-1] if v else None),
    "abs": ((1,), abs), "coalesce": (tuple(range(1, 33)), None),
    "exists": ((1,), lambda v: True),                # exists(n.p): the null case is answered below
    "type": ((1,), lambda r: r.type), "id": ((1,), lambda x: x.id), "elementid": ((1,), lambda x: x.element_id),
    "labels": ((1,), lambda n: sorted(n.labels)), "keys": ((1,), lambda x: sorted(x.keys())),
    "properties": ((1,), lambda x: dict(x.items())),
    "startnode": ((1,), lambda r: r.start_node), "endnode": ((1,), lambda r: r.end_node),
    "nodes": ((1,), lambda p: list(p.nodes)), "relationships": ((1,), lambda p: list(p.relationships)),
}


def conjuncts(e):
    if e is None:
        return []
    if e[0] == "and":
        return conjuncts(e[1]) + conjuncts(e[2])
    return [e]


# ------------------------------------------------------------------------------------------ plans
@dataclass
class NodeScanPlan:
    var: str
    label: str | None
    where: object
    id_equals: str | None          # fast path: WHERE var.id = <string>
    ret: Clause


@dataclass
class VarLenPlan:
    path_var: str | None
    node_vars: list                # variable names of the pattern nodes (fixed-length patterns) or [a, b]
    rel_vars: list
    direction: str                 # out | in | und
    min_hops: int
    max_hops: int
    kind_prop: str                 # the property the WHERE clause tests ('kind')
    src_kind: str | None
    dest_kind: str | None
    pos_kinds: list | None         # fixed-length patterns: required kind per position (None = any)
    excl_kinds: list
    inter_kinds: list
    node_simple: bool
    ret: Clause


@dataclass
class ChainStep:
    rel_var: str | None
    rel_types: list
    key: object                    # literal value of r.key, or None
    start_var: str                 # as stored: start -> end (undirected: as written, either way round may match)
    end_var: str
    undirected: bool = False


@dataclass
class ChainPlan:
    seed_var: str
    seed_where: object             # predicate over the seed alone (AST) or None
    seed_limit: int | None
    labels: dict                   # variable -> label (or None)
    steps: list                    # in walk order
    walk: list                     # per step: (from_var, to_var, 'out' | 'in')
    ret: Clause
    unique: list = field(default_factory=list)   # groups of step indices written in ONE MATCH clause: Cypher binds
                                                 # no relationship twice within a clause (across clauses it may)
    residual: list = field(default_factory=list) # predicates evaluated on the finished rows (host side)
    hints: dict = field(default_factory=dict)    # variable -> [(property, value)]: equality tests that may name its kind


@dataclass
class StatePlan:
    entity_var: str
    entity_label: str
    state_var: str
    state_label: str | None
    entity_id: str
    lo: str                        # r.tmin <= lo
    hi: str                        # r.tmax >  hi
    ret: Clause
    late: tuple = ()               # the expressions entity_id, lo and hi came from (a plan kept per skeleton is bound
                                   # to each statement's own values: bound())

    def bound(self, params):
        e, lo, hi = self.late
        return StatePlan(self.entity_var, self.entity_label, self.state_var, self.state_label, _late(e, params),
                         _late(lo, params), _late(hi, params), self.ret, self.late)


def _lit(e, params):
    """Constant value of a literal / parameter expression, else raise."""
    if e[0] == "lit":
        return e[1]
    if e[0] == "param":
        if e[1] not in params:
            raise CypherSyntaxError(f"Expected parameter(s): {e[1]}", "Neo.ClientError.Statement.ParameterMissing")
        noted = getattr(_tls, "noted", None)
        if noted is not None:
            noted.append((e[1], params[e[1]]))
        return params[e[1]]
    if e[0] == "list":
        return [_lit(x, params) for x in e[1]]
    raise CypherSyntaxError("Only literals and parameters are supported on the right-hand side of this predicate")


def _late(e, params):
    """_lit for a value that is bound per statement: it does not make the plan specific to the value."""
    keep, _tls.noted = getattr(_tls, "noted", None), None
    try:
        return _lit(e, params)
    finally:
        _tls.noted = keep


def _is_prop_of(e, var=None):
    return e[0] == "prop" and e[1][0] == "var" and (var is None or e[1][1] == var)


def _unsupported(what):
    return CypherSyntaxError(f"This executor does not support {what}; supported shapes are listed in "
                             f"k8s-llm-rca_b200/cypher.py")


def compile_query(text, parameters=None, bind=None):
    """Parse `text` and recognise its shape.  Returns one of the *Plan dataclasses.
    bind: a dict that receives the string literals lifted out of the text (see parse_shared); a caller that
    evaluates the plan's expressions later must pass the dict it evaluates with."""
    params = dict(parameters or {})
    if bind is None:
        return _compile_clauses(parse(text), params)
    clauses, lifted, skel = _parse_shared(text)
    params.update(lifted)
    bind.update(lifted)
    if skel is None:
        return _compile_clauses(clauses, params)
    variants = _PLANS.setdefault(skel, [])
    absent = _PLANS                                  # a value no parameter has
    for looked_at, plan in variants:
        try:
            same = all(params.get(n, absent) == v for n, v in looked_at)
        except (TypeError, ValueError):              # a parameter value that does not compare to a plain truth value
            same = False
        if same:
            return plan.bound(params) if isinstance(plan, StatePlan) else plan
    _tls.noted = looked_at = []
    try:
        plan = _compile_clauses(clauses, params)
    finally:
        _tls.noted = None
    if len(variants) < 64:
        variants.append((looked_at, plan))
    return plan


def _compile_clauses(clauses, params):
    matches = [c for c in clauses if c.kind == "MATCH"]
    ret = clauses[-1]
    if not matches:
        raise _unsupported("queries without MATCH")
    pats = [c.patterns[0] for c in matches]

    if len(matches) == 1 and len(matches[0].patterns) == 1:
        p, where = pats[0], matches[0].where
        if not p.rels:
            return _compile_node_scan(p, where, ret, params)
        if any(r.varlen for r in p.rels):
            return _compile_varlen(p, where, ret, params)
        if len(p.rels) == 1 and p.rels[0].types == ["HasState"] and _mentions(where, p.rels[0].var, ("tmin", "tmax")):
            return _compile_state(p, where, ret, params)
        if all(not r.types and not r.props for r in p.rels):
            return _compile_fixed_untyped(p, where, ret, params)
    return _compile_chain(clauses, params)


def _mentions(where, var, props):
    return var is not None and any(
        c[0] == "cmp" and _is_prop_of(c[2], var) and c[2][2] in props for c in conjuncts(where))


def _compile_node_scan(p, where, ret, params):
    n = p.nodes[0]
    if n.var is None:
        raise _unsupported("anonymous node scans")
    if len(n.labels) > 1:
        raise _unsupported("multi-label node patterns")
    w = where
    for k, v in n.props.items():
        eq = ("cmp", "=", ("prop", ("var", n.var), k), v)
        w = eq if w is None else ("and", w, eq)
    id_eq = None
    cj = conjuncts(w)
    if len(cj) == 1 and cj[0][0] == "cmp" and cj[0][1] == "=" and _is_prop_of(cj[0][2], n.var) and cj[0][2][2] == "id":
        v = _lit(cj[0][3], params)
        if isinstance(v, str):
            id_eq = v
    return NodeScanPlan(n.var, n.labels[0] if n.labels else None, w, id_eq, ret)


def _compile_varlen(p, where, ret, params):
    if len(p.rels) != 1:
        raise _unsupported("variable-length relationships mixed with other hops")
    r, a, b = p.rels[0], p.nodes[0], p.nodes[1]
    if r.types or r.props or a.labels or b.labels or a.props or b.props:
        raise _unsupported("typed or labelled variable-length patterns")
    lo, hi = r.varlen
    if hi is None:
        raise _unsupported("unbounded variable-length patterns")
    if lo < 1 or hi < lo:
        raise CypherSyntaxError(f"Invalid variable-length range *{lo}..{hi}")
    plan = VarLenPlan(p.path_var, [a.var, b.var], [r.var], r.direction, lo, hi, "kind", None, None, None, [], [],
                      False, ret)
    for c in conjuncts(where):
        if c[0] == "cmp" and c[1] == "=" and _is_prop_of(c[2]) and c[2][1][1] in (a.var, b.var):
            plan.kind_prop = c[2][2]
            val = _lit(c[3], params)
            if c[2][1][1] == a.var:
                plan.src_kind = val
            else:
                plan.dest_kind = val
        elif c[0] == "quant" and c[1] == "all" and _is_nodes_of(c[3], p.path_var):
            pred, v = c[4], c[2]
            if pred[0] == "quant" and pred[1] == "single" and _is_nodes_of(pred[3], p.path_var):
                inner = pred[4]                  # single(x in nodes(path) WHERE x = node)
                if inner[0] == "cmp" and inner[1] == "=" and {inner[2], inner[3]} == {("var", v), ("var", pred[2])}:
                    plan.node_simple = True
                    continue
                raise _unsupported("this single(...) predicate")
            if pred[0] == "not" and pred[1][0] == "cmp" and pred[1][1] == "in" and _is_prop_of(pred[1][2], v):
                plan.kind_prop = pred[1][2][2]
                plan.excl_kinds = list(_lit(pred[1][3], params))
                continue
            raise _unsupported("this all(...) predicate")
        elif c[0] == "or":
            inter = _inter_kinds(c, p.path_var, params)
            if inter is None:
                raise _unsupported("this OR predicate in a variable-length query")
            plan.inter_kinds = inter
        elif c[0] == "quant" and c[1] == "any":
            inter = _inter_kinds(c, p.path_var, params)
            if inter is None:
                raise _unsupported("this any(...) predicate")
            plan.inter_kinds = inter
        else:
            raise _unsupported("this predicate in a variable-length query")
    if plan.src_kind is None:
        raise _unsupported("variable-length queries without a kind test on the start node")
    return plan


def _is_nodes_of(e, path_var):
    return e[0] == "call" and e[1] == "nodes" and len(e[2]) == 1 and e[2][0] == ("var", path_var)


def _inter_kinds(c, path_var, params):
    """($p IS NULL OR size($p) = 0 OR any(node in nodes(path)[1..-1] WHERE node.kind in $p)) -> list of kinds
    ([] when the parameter is null or empty: the rule is then vacuous)."""
    parts = []

    def flat(e):
        if e[0] == "or":
            flat(e[1])
            flat(e[2])
        else:
            parts.append(e)
    flat(c)
    anyq = [x for x in parts if x[0] == "quant" and x[1] == "any"]
    if len(anyq) != 1:
        return None
    q = anyq[0]
    lst = q[3]
    if not (lst[0] == "slice" and _is_nodes_of(lst[1], path_var) and lst[2] == ("lit", 1) and lst[3] == ("neg", ("lit", 1))):
        return None
    pred = q[4]
    if not (pred[0] == "cmp" and pred[1] == "in" and _is_prop_of(pred[2], q[2])):
        return None
    kinds = _lit(pred[3], params)
    for x in parts:
        if x is q:
            continue
        ok = (x[0] == "isnull" and not x[2] and x[1] == pred[3]) or \
             (x[0] == "cmp" and x[1] == "=" and x[2] == ("call", "size", [pred[3]]) and x[3] == ("lit", 0))
        if not ok:
            return None
    return list(kinds) if kinds else []


def _compile_fixed_untyped(p, where, ret, params):
    dirs = {r.direction for r in p.rels}
    if len(dirs) != 1:
        raise _unsupported("fixed patterns that mix relationship directions")
    n = len(p.rels)
    pos = [None] * (n + 1)
    kind_prop = "kind"
    names = [x.var for x in p.nodes]
    for c in conjuncts(where):
        if c[0] == "cmp" and c[1] == "=" and _is_prop_of(c[2]) and c[2][1][1] in names:
            kind_prop = c[2][2]
            pos[names.index(c[2][1][1])] = _lit(c[3], params)
        else:
            raise _unsupported("this predicate in an untyped fixed-length pattern")
    if pos[0] is None:
        raise _unsupported("untyped patterns without a kind test on the first node")
    return VarLenPlan(p.path_var, names, [r.var for r in p.rels], dirs.pop(), n, n, kind_prop, pos[0], pos[-1], pos,
                      [], [], False, ret)


def _compile_state(p, where, ret, params):
    a, r, b = p.nodes[0], p.rels[0], p.nodes[1]
    if r.direction != "out" or not a.labels:
        raise _unsupported("HasState patterns that are not (entity:Kind)-[:HasState]->(state)")
    ent_id = lo = hi = None
    for c in conjuncts(where):
        if c[0] != "cmp":
            raise _unsupported("this predicate in a state query")
        if _is_prop_of(c[2], a.var) and c[2][2] == "id" and c[1] == "=":
            ent_id = c[3]
        elif _is_prop_of(c[2], r.var) and c[2][2] == "tmin" and c[1] == "<=":
            lo = c[3]
        elif _is_prop_of(c[2], r.var) and c[2][2] == "tmax" and c[1] == ">":
            hi = c[3]
        else:
            raise _unsupported("this predicate in a state query")
    if ent_id is None or lo is None or hi is None:
        raise _unsupported("state queries without id, tmin <= and tmax > predicates")
    return StatePlan(a.var, a.labels[0], b.var, b.labels[0] if b.labels else None, _late(ent_id, params),
                     _late(lo, params), _late(hi, params), ret, (ent_id, lo, hi))


def _compile_chain(clauses, params):
    labels, hops, preds, seed_limit, limited_vars, limit_at = {}, [], [], None, None, None
    first_vars = None
    anon = [0]

    def note(n, ci):
        if n.var is None:                            # '()' or '(:Pod)': a hidden variable (RETURN * does not list it)
            n.var = f" anon{anon[0]}"
            anon[0] += 1
        if len(n.labels) > 1:
            raise _unsupported("multi-label node patterns")
        if n.labels:
            if labels.get(n.var) not in (None, n.labels[0]):
                raise CypherSyntaxError(f"Variable `{n.var}` is given two different labels")
            labels[n.var] = n.labels[0]
        labels.setdefault(n.var, None)
        for k, v in n.props.items():
            preds.append((ci, ("cmp", "=", ("prop", ("var", n.var), k), v)))

    clause_of = []                                   # per hop: index of the MATCH clause it was written in
    for ci, c in enumerate(clauses[:-1]):
        if c.kind == "MATCH":
            for p in c.patterns:                     # 'MATCH p1, p2' = both patterns in one clause
                if p.path_var:
                    raise _unsupported("path variables over typed patterns")
                for n in p.nodes:
                    note(n, ci)
                if first_vars is None:
                    first_vars = [n.var for n in p.nodes]
                for i, r in enumerate(p.rels):
                    if r.varlen:
                        raise _unsupported("variable-length hops inside typed patterns")
                    if not r.types:
                        raise _unsupported("untyped hops mixed with typed ones")
                    a, b = p.nodes[i].var, p.nodes[i + 1].var
                    start, end = (b, a) if r.direction == "in" else (a, b)
                    hops.append(ChainStep(r.var, list(r.types), r.props.get("key"), start, end, r.direction == "und"))
                    clause_of.append(ci)
                    for k, v in r.props.items():
                        if k != "key":
                            if r.var is None:
                                raise _unsupported("property maps other than {key: ...} on an unnamed relationship")
                            preds.append((ci, ("cmp", "=", ("prop", ("var", r.var), k), v)))
            preds.extend((ci, x) for x in conjuncts(c.where))
        elif c.kind == "WITH":
            if c.limit is not None:
                if seed_limit is not None:
                    raise _unsupported("more than one WITH ... LIMIT")
                seed_limit, limit_at = c.limit, ci
                limited_vars = [it[0][1] for it in c.items if it[0][0] == "var"]
            for it in c.items:
                if it[0] == ("star",):
                    continue
                if it[0][0] != "var" or (it[1] and it[1] != it[0][1]):
                    raise _unsupported("WITH items other than plain variables")
    if not hops:
        raise _unsupported("this query shape")
    used = {h.start_var for h in hops} | {h.end_var for h in hops}
    if any(v not in used for v in labels):
        raise _unsupported("patterns that are not connected to the others (cartesian products)")

    # Predicates: `r.key = <literal>` selects the step's edge types; predicates over the seed alone written before
    # its LIMIT choose the seeds; everything else is evaluated on the finished rows (`residual`).
    rel_vars = {h.rel_var: h for h in hops if h.rel_var}
    single, residual = [], []
    for ci, c in preds:
        if c[0] == "cmp" and c[1] == "=" and _is_prop_of(c[2]) and c[2][1][1] in rel_vars and c[2][2] == "key" \
                and c[3][0] in ("lit", "param") and rel_vars[c[2][1][1]].key is None:
            rel_vars[c[2][1][1]].key = c[3]
            continue
        vars_ = _vars_of(c)
        unknown = [v for v in vars_ if v not in labels and v not in rel_vars]
        if unknown:
            raise CypherSyntaxError(f"Variable `{unknown[0]}` not defined")
        if len(vars_) == 1 and next(iter(vars_)) in labels:
            single.append((ci, next(iter(vars_)), c))
        else:
            residual.append(c)
    if limited_vars:
        seed_var = limited_vars[0]
    else:
        def selective(c):                            # the lookups the engine answers itself (message arena, id table)
            return c[0] == "cmp" and _is_prop_of(c[2]) and ((c[1] == "contains" and c[2][2] == "message") or
                                                            (c[1] == "=" and c[2][2] == "id"))
        # ... else an END of the line of hops (the engine walks a line from one end), one with a predicate if any
        degree = {}
        for h in hops:
            for v in (h.start_var, h.end_var):
                degree[v] = degree.get(v, 0) + 1
        ends = [v for v in labels if degree.get(v) == 1]
        seed_var = next((v for _, v, c in single if selective(c)), None) or \
            next((v for _, v, _ in single if v in ends), None) or \
            (first_vars[0] if first_vars[0] in ends or not ends else ends[0])
    if seed_var not in labels:
        raise CypherSyntaxError(f"Variable `{seed_var}` not defined")
    seed_where, hints = None, {}
    for ci, v, c in single:
        if v == seed_var and (limit_at is None or ci < limit_at):
            seed_where = c if seed_where is None else ("and", seed_where, c)
            continue
        residual.append(c)
        if c[0] == "cmp" and c[1] == "=" and _is_prop_of(c[2], v) and c[3][0] in ("lit", "param"):
            hints.setdefault(v, []).append((c[2][2], _lit(c[3], params)))   # `pod.kind2 = 'Pod'` names the kind
    for h in hops:
        if h.key is not None:
            h.key = _lit(h.key, params)

    # order the hops as one walk from the seed
    bound, walk, steps, todo = {seed_var}, [], [], list(hops)
    cur = seed_var
    while todo:
        nxt = [h for h in todo if h.start_var in bound or h.end_var in bound]
        if not nxt:
            raise _unsupported("patterns that are not connected to the filtered node")
        h = next((x for x in nxt if cur in (x.start_var, x.end_var)), None)
        if h is None:
            raise _unsupported("branching patterns (every hop must extend the previous one)")
        todo.remove(h)
        if h.start_var in bound and h.end_var in bound:
            raise _unsupported("patterns that close a cycle")
        if h.start_var == cur:
            walk.append((h.start_var, h.end_var, "und" if h.undirected else "out"))
            cur = h.end_var
        else:
            walk.append((h.end_var, h.start_var, "und" if h.undirected else "in"))
            cur = h.start_var
        bound.add(cur)
        steps.append(h)
    groups = {}
    for i, h in enumerate(steps):
        groups.setdefault(clause_of[next(j for j, x in enumerate(hops) if x is h)], []).append(i)
    unique = [g for g in groups.values() if len(g) > 1]
    return ChainPlan(seed_var, seed_where, seed_limit, labels, steps, walk, clauses[-1], unique, residual, hints)


def _vars_of(e):
    out = set()

    def rec(x):
        if isinstance(x, tuple):
            if x and x[0] == "var":
                out.add(x[1])
            for y in x[1:]:
                rec(y)
        elif isinstance(x, list):
            for y in x:
                rec(y)
    rec(e)
    return out
