"""jsfxc.py -- DSP-JSFX -> CUDA translator for the catalogue's self-contained leaves.

Build-time tool.  Reads a .jsfx script and writes one .cu translation unit holding the script's four
sections (@init / @slider / @block / @sample) and its user functions as device code for ONE stream per
thread (register / local-memory resident variables, mem[] as a stream-minor structure-of-arrays heap in
HBM), plus the descriptor the generic plugin back end (csrc/p_generic.cu) registers.  The product never
runs this at run time; generated files are committed under csrc/gen/ so the GPU box only compiles them.

Semantics restated from the reference's AOT compiler (file:line in /root/reference/dsp_jsfx_aot.py):
  * tokens / grammar: newline and ';' separate statements, a newline followed by an infix operator
    continues the expression (:594-630); precedence table :370-389; prefix + - ! bind tighter than any
    binary operator (:680-689); `c ? a` without ':' yields 0 (:658-675); ( a; b ) sequences (:790-819);
    loop(n, body) (:703-739); while (c) body and if (c) a else b keywords (:473-498);
    function f(a b) local(x) ( body ) (:500-577)
  * expressions (:4263-5590): all values double; truthiness is `!= 0` exactly (:3725); compares yield
    1.0 / 0.0 ordered; | & << >> % work on trunc-to-int64 wrapped to int32 (:4107-4110, :4355-4381), shift
    count & 31, >> arithmetic; ^ is pow(); unary minus is 0 - x; min / max are olt / ogt selects
    (:5223-5232); sqr, sign (:5234-5247); invsqrt keeps the float bit trick (:5256-5277); rand() is a
    per-instance MT19937 seeded 0x4141F00D (:3880-4060, :5294-5323); gfx_* / sprintf / str* evaluate
    their arguments and return 0 (:5200-5212); && || short-circuit and return 1.0 / 0.0 (:5592-5633)
  * a[b] = mem[trunc(a + b + 1e-5)], negative -> 0 (:4062-4103); assignment evaluates the right-hand side
    first, then the target address (:4386-4492); memset returns dest (:5439-5495)
  * loop count: trunc toward 0, negative -> 0, evaluated once (:5663-5707)
  * user functions are specialised per calling SECTION; local() variables are persistent state named by
    (section, function, local) (:1844-1848, :1905-2048); parameters are by value
  * instance() / `this.` pseudo-objects (:1905-2048): a function that declares instance() variables or mentions `this` is
    specialised per call-site namespace as well (`lp.f(x)`: instance s1 -> the global `lp.s1`; called bare, the function
    is its own namespace); local() storage stays per (section, function)
  * fft / ifft / fft_real / ifft_real / fft_permute / fft_ipermute / convolve_c (:5498-5564): the runtime hooks become
    per-stream device transforms on the heap with WDL's ordering and scaling contract (csrc/jsfx_rt.cuh)
Not supported (translation fails loudly): gmem, midi, file and message builtins, string slider files.
"""
from __future__ import annotations

import re
import sys
from pathlib import Path

# ------------------------------------------------------------------------------------------------------
# lexer
# ------------------------------------------------------------------------------------------------------
OPS3 = ()
OPS2 = ("+=", "-=", "*=", "/=", "%=", "^=", "|=", "&=", "~=", "==", "!=", "<=", ">=", "&&", "||", "<<", ">>")
OPS1 = "+-*/%^|&<>=!?:~"
KEYWORDS = ("if", "else", "while")


class Tok:
    __slots__ = ("kind", "text", "line")

    def __init__(self, kind, text, line):
        self.kind, self.text, self.line = kind, text, line

    def __repr__(self):
        return f"{self.kind}:{self.text!r}@{self.line}"


def lex(src: str, line0: int = 1):
    toks, i, n, line = [], 0, len(src), line0
    while i < n:
        c = src[i]
        if c == "\n":
            toks.append(Tok("eol", "\n", line)); line += 1; i += 1
        elif c in " \t\r":
            i += 1
        elif src.startswith("//", i):
            while i < n and src[i] != "\n":
                i += 1
        elif src.startswith("/*", i):
            j = src.find("*/", i + 2)
            j = n if j < 0 else j + 2
            line += src.count("\n", i, j); i = j
        elif c.isdigit() or (c == "." and i + 1 < n and src[i + 1].isdigit()):
            m = re.match(r"0[xX][0-9a-fA-F]+|(?:\d+\.?\d*|\.\d+)(?:[eE][-+]?\d+)?", src[i:])
            toks.append(Tok("num", m.group(0), line)); i += len(m.group(0))
        elif c.isalpha() or c in "_#$":
            if c == "$" and i + 1 < n and src[i + 1] == "'":          # $'c' character constant
                toks.append(Tok("num", str(ord(src[i + 2])), line)); i += 4
                continue
            m = re.match(r"[A-Za-z_#$][A-Za-z0-9_.#$]*", src[i:])
            t = m.group(0)
            toks.append(Tok("kw" if t in KEYWORDS else "ident", t, line)); i += len(t)
        elif c == '"' or c == "'":
            j = i + 1
            while j < n and src[j] != c:
                j += 2 if src[j] == "\\" else 1
            toks.append(Tok("str", src[i + 1:j], line)); line += src.count("\n", i, j); i = j + 1
        elif src[i:i + 2] in OPS2:
            toks.append(Tok("op", src[i:i + 2], line)); i += 2
        elif c in OPS1:
            toks.append(Tok("op", c, line)); i += 1
        elif c in "()[],":
            toks.append(Tok("punc", c, line)); i += 1
        elif c == ";":
            toks.append(Tok("semi", ";", line)); i += 1
        else:
            raise SyntaxError(f"line {line}: unexpected character {c!r}")
    toks.append(Tok("eof", "", line))
    return toks


# ------------------------------------------------------------------------------------------------------
# parser (nodes are tuples: (kind, ...))
# ------------------------------------------------------------------------------------------------------
PREC = {"=": 1, "+=": 1, "-=": 1, "*=": 1, "/=": 1, "%=": 1, "^=": 1, "|=": 1, "&=": 1, "~=": 1,
        "||": 3, "|": 3, "&&": 4, "==": 5, "!=": 5, "&": 5, "<": 6, "<=": 6, ">": 6, ">=": 6, "<<": 6, ">>": 6,
        "+": 7, "-": 7, "*": 8, "/": 8, "%": 8, "^": 9}
ASSIGN = {"=", "+=", "-=", "*=", "/=", "%=", "^=", "|=", "&=", "~="}
TERN = 2


class Parser:
    def __init__(self, src, line0=1):
        self.t = lex(src, line0)
        self.i = 0

    @property
    def cur(self):
        return self.t[self.i]

    @property
    def nxt(self):
        return self.t[min(self.i + 1, len(self.t) - 1)]

    def adv(self):
        self.i += 1

    def err(self, msg):
        raise SyntaxError(f"line {self.cur.line}: {msg} (at {self.cur!r})")

    def eat(self, kind, text=None):
        c = self.cur
        if c.kind != kind or (text is not None and c.text != text):
            self.err(f"expected {text or kind}")
        self.adv()
        return c

    def is_p(self, ch):
        return self.cur.kind == "punc" and self.cur.text == ch

    def skip_seps(self):
        while self.cur.kind in ("eol", "semi"):
            self.adv()

    def skip_eol(self):
        while self.cur.kind == "eol":
            self.adv()

    def program(self):
        out = []
        self.skip_seps()
        while self.cur.kind != "eof":
            out.append(self.stmt())
            self.skip_seps()
        return out

    def stmt(self):
        if self.cur.kind == "kw" and self.cur.text == "if":
            return self.p_if()
        if self.cur.kind == "kw" and self.cur.text == "while":
            return self.p_while()
        if self.cur.kind == "ident" and self.cur.text == "function":
            return self.p_function()
        return self.expr(0)

    def p_if(self):
        self.eat("kw", "if"); self.eat("punc", "(")
        c = self.expr(0)
        self.skip_seps(); self.eat("punc", ")"); self.skip_seps()
        a = self.expr(0); self.skip_seps()
        b = None
        if self.cur.kind == "kw" and self.cur.text == "else":
            self.adv(); self.skip_seps()
            b = self.expr(0); self.skip_seps()
        return ("if", c, a, b)

    def p_while(self):
        self.eat("kw", "while"); self.eat("punc", "(")
        c = self.expr(0)
        self.skip_seps(); self.eat("punc", ")"); self.skip_seps()
        return ("while", c, self.expr(0))

    def name_list(self):
        out = []
        self.eat("punc", "("); self.skip_seps()
        while not self.is_p(")"):
            out.append(self.eat("ident").text)
            self.skip_seps()
            if self.is_p(","):
                self.adv(); self.skip_seps()
        self.eat("punc", ")")
        return out

    def p_function(self):
        self.eat("ident", "function")
        name = self.eat("ident").text
        params = self.name_list()
        locals_, instances = [], []
        self.skip_seps()
        while self.cur.kind == "ident" and self.cur.text in ("local", "instance", "global", "static"):
            q = self.cur.text
            self.adv()
            names = self.name_list()
            if q == "local":
                locals_ += names
            elif q == "instance":
                instances += names
            self.skip_seps()
        if not self.is_p("("):
            self.err("expected '(' to start function body")
        body = self.primary()
        return ("fundef", name, params, locals_, instances, body)

    def cont_op(self, tok, min_prec):
        if tok.kind != "op":
            return False
        if tok.text == "?":
            return TERN >= min_prec
        if tok.text in (":", "+", "-", "!"):
            return False
        p = PREC.get(tok.text)
        return p is not None and p >= min_prec

    def expr(self, min_prec):
        lhs = self.prefix()
        while True:
            while self.cur.kind == "eol" and (self.nxt.kind == "eol" or self.cont_op(self.nxt, min_prec)):
                self.adv()
            if self.cur.kind != "op":
                break
            op = self.cur.text
            if op in ("?", ":"):
                break
            p = PREC.get(op)
            if p is None or p < min_prec:
                break
            self.adv()
            rhs = self.expr(p if op in ASSIGN else p + 1)
            if op in ASSIGN:
                if lhs[0] not in ("var", "index") and not (lhs[0] == "call" and lhs[1] in ("slider", "spl")):
                    self.err("bad assignment target")
                lhs = ("assign", op, lhs, rhs)
            else:
                lhs = ("bin", op, lhs, rhs)
        while self.cur.kind == "eol" and (self.nxt.kind == "eol" or (self.nxt.kind == "op" and self.nxt.text == "?")):
            self.adv()
        if self.cur.kind == "op" and self.cur.text == "?" and TERN >= min_prec:
            self.adv(); self.skip_seps()
            a = self.expr(0); self.skip_seps()
            if self.cur.kind == "op" and self.cur.text == ":":
                self.adv(); self.skip_seps()
                b = self.expr(0)
            else:
                b = ("num", 0.0)
            lhs = ("tern", lhs, a, b)
        return lhs

    def prefix(self):
        self.skip_eol()
        if self.cur.kind == "op" and self.cur.text in ("+", "-", "!"):
            op = self.cur.text
            self.adv()
            return ("un", op, self.prefix())
        return self.postfix()

    def seq_item(self):
        if self.cur.kind == "kw" and self.cur.text == "if":
            return self.p_if()
        if self.cur.kind == "kw" and self.cur.text == "while":
            return self.p_while()
        return self.expr(0)

    def postfix(self):
        node = self.primary()
        while True:
            if self.is_p("("):
                if node[0] != "var":
                    self.err("can only call a named function")
                fn = node[1]
                self.adv()
                if fn == "loop":
                    self.skip_seps()
                    count = self.expr(0); self.skip_seps()
                    if self.is_p(","):
                        self.adv()
                    self.skip_seps()
                    items = []
                    while not self.is_p(")"):
                        items.append(self.seq_item()); self.skip_seps()
                    self.adv()
                    node = ("loop", count, ("seq", items) if len(items) != 1 else items[0])
                    continue
                args = []
                self.skip_seps()
                while not self.is_p(")"):
                    args.append(self.expr(0)); self.skip_seps()
                    if self.is_p(","):
                        self.adv(); self.skip_seps()
                self.adv()
                node = ("call", fn, args)
                continue
            if self.is_p("["):
                self.adv(); self.skip_seps()
                idx = ("num", 0.0) if self.is_p("]") else self.expr(0)
                self.skip_seps(); self.eat("punc", "]")
                node = ("index", node, idx)
                continue
            return node

    def primary(self):
        c = self.cur
        if c.kind == "num":
            self.adv()
            return ("num", float(int(c.text, 16)) if c.text[:2].lower() == "0x" else float(c.text))
        if c.kind == "str":
            self.adv()
            return ("str", c.text)
        if c.kind == "ident":
            self.adv()
            return ("var", c.text)
        if self.is_p("("):
            self.adv(); self.skip_seps()
            items = []
            while not self.is_p(")"):
                items.append(self.seq_item()); self.skip_seps()
            self.adv()
            return items[0] if len(items) == 1 else ("seq", items)
        self.err("expected number, identifier or '('")


# ------------------------------------------------------------------------------------------------------
# script = header + sections
# ------------------------------------------------------------------------------------------------------
SECTIONS = ("init", "slider", "block", "sample")


def split_script(text: str):
    header, secs, cur, start = [], {}, None, {}
    for ln, line in enumerate(text.splitlines(), 1):
        m = re.match(r"^@(\w+)", line)
        if m:
            cur = m.group(1)
            secs.setdefault(cur, []); start.setdefault(cur, ln + 1)
            continue
        (header if cur is None else secs[cur]).append(line)
    return "\n".join(header), {k: ("\n".join(v), start[k]) for k, v in secs.items()}


def pin_counts(header: str, reads: set, writes: set):
    """Channel counts as the reference infers them (dsp_jsfx_aot.py parse_pin_hints :1110-1136, infer_spl_io :1662):
    in_pin / out_pin lines win ('none' = 0 pins); else a read of splN needs N+1 inputs, a write N+1 outputs; an
    undeclared empty side mirrors the other; nothing at all = stereo."""
    saw, cnt = {"in_pin": False, "out_pin": False}, {"in_pin": 0, "out_pin": 0}
    for raw in header.splitlines():
        line = raw.split("//", 1)[0].split(";", 1)[0]
        m = re.match(r"^\s*(in_pin|out_pin)\s*:\s*(.*?)\s*$", line, re.I)
        if not m:
            continue
        k = m.group(1).lower()
        saw[k] = True
        if m.group(2).strip().lower() == "none":
            cnt[k] = 0
        else:
            cnt[k] += 1
    d_in = cnt["in_pin"] if saw["in_pin"] else None
    d_out = cnt["out_pin"] if saw["out_pin"] else None
    n_in = (max(reads) + 1) if reads else 0
    n_out = (max(writes) + 1) if writes else 0
    if d_in is not None:
        n_in = d_in
    if d_out is not None:
        n_out = d_out
    if d_in is None and d_out is None and n_in == 0 and n_out == 0:
        n_in = n_out = 2
    if d_in is None and n_in == 0 and n_out > 0:
        n_in = n_out
    if d_out is None and n_out == 0 and n_in > 0:
        n_out = n_in
    return min(n_in, 64), min(n_out, 64)


MATH1 = {"sin", "cos", "tan", "asin", "acos", "atan", "exp", "log", "log10", "sqrt", "floor", "ceil", "abs"}
MATH2 = {"pow", "atan2"}
NOOP_CALLS = {"sprintf", "printf", "strcpy", "strcat", "strcmp", "strlen", "str_getchar", "str_setchar", "str_insert",
              "str_delete", "str_mid", "strncpy", "file_read", "file_write", "file_string"}
FFT_CALLS = {"fft": "c.fft({0}, {1}, false)", "ifft": "c.fft({0}, {1}, true)", "fft_real": "c.fft_real({0}, {1}, false)",
             "ifft_real": "c.fft_real({0}, {1}, true)", "fft_permute": "c.fft_permute({0}, {1}, false)",
             "fft_ipermute": "c.fft_permute({0}, {1}, true)"}
UNSUPPORTED = {"slider_next_chg", "midisend", "midirecv", "midisend_buf", "midirecv_buf",
               "file_open", "file_close", "file_mem", "file_riff", "file_avail"}
STATE_BUILTINS = {"srate": "srate", "samplesblock": "samplesblock"}


def cname(name: str) -> str:
    return "v_" + re.sub(r"[^A-Za-z0-9_]", lambda m: "_%02x" % ord(m.group(0)), name)


def uses_this(n) -> bool:
    """Does a parsed node mention `this` / `this.x` (as a variable or as a call prefix)?  (_node_uses_relative_namespace, :1860-1898)"""
    if isinstance(n, tuple):
        if n and n[0] in ("var", "call") and isinstance(n[1], str) and (n[1] == "this" or n[1].startswith("this.")):
            return True
        return any(uses_this(x) for x in n[1:])
    if isinstance(n, list):
        return any(uses_this(x) for x in n)
    return False


class Unit:
    """One script: variable table, function specialisations, emitted section bodies."""

    def __init__(self, key: str, text: str):
        self.key = key
        header, secs = split_script(text)
        self.header = header
        self.slider_lines = [l for l in header.splitlines() if re.match(r"^\s*slider\d+\s*:", l)]
        self.fundefs = {}
        self.progs = {}
        for s in SECTIONS:
            src, l0 = secs.get(s, ("", 1))
            prog = Parser(src, l0).program()
            keep = []
            for n in prog:
                if n[0] == "fundef":
                    self.fundefs[n[1]] = n
                else:
                    keep.append(n)
            self.progs[s] = keep
        # a function is specialised per namespace when it declares instance() variables or mentions `this` (:1988-1991)
        self.needs_ns = {name: bool(fd[4]) or uses_this(fd[5]) for name, fd in self.fundefs.items()}
        self.uses_fft = False
        self.vars = {}            # jsfx name -> slot
        self.used_spl = set()
        self.spl_reads, self.spl_writes = set(), set()
        self.used_sliders = set()
        self.spec = {}            # (section, fn) -> C name
        self.spec_code = []       # emitted function definitions, callee first
        self.in_progress = set()
        self.fn_trace = {}        # specialised function -> access trace of its body (stage planner)
        self.body, self.trace, self.stmt_lines = {}, {}, {}
        for s in SECTIONS:
            em = Emitter(self, s, {}, set())
            self.body[s] = em.run(self.progs[s])
            self.trace[s], self.stmt_lines[s] = em.trace, em.stmt_lines
        self.n_in, self.n_out = pin_counts(header, self.spl_reads, self.spl_writes)
        self.alias = {}           # slider index0 -> var name
        for l in self.slider_lines:
            m = re.match(r"^\s*slider(\d+)\s*:\s*([A-Za-z_][A-Za-z0-9_.]*)\s*=", l)
            if m:
                self.alias[int(m.group(1)) - 1] = m.group(2)
                self.slot(m.group(2))

    def slot(self, name: str) -> int:
        if name not in self.vars:
            self.vars[name] = len(self.vars)
        return self.vars[name]

    def resolve_call(self, fn: str, current_ns):
        """User-function call name -> (base function, call namespace) as lower_user_functions does (:2017-2027): `f(...)` is a
        plain call; `a.b.f(...)` calls f with namespace a.b, where a leading `this` is the caller's own namespace."""
        if fn in self.fundefs:
            return fn, None
        parts = fn.split(".")
        if len(parts) >= 2 and parts[-1] in self.fundefs:
            prefix = ".".join(parts[:-1])
            if prefix == "this":
                ns = current_ns
            elif prefix.startswith("this."):
                suffix = prefix[5:]
                ns = (f"{current_ns}.{suffix}" if suffix else current_ns) if current_ns else (suffix or current_ns)
            else:
                ns = prefix
            return parts[-1], ns
        return None, None

    def specialise(self, section: str, fn: str, call_ns=None) -> str:
        # per caller section, and per namespace where the function needs one; without a call namespace such a function is
        # its own namespace (:2029-2040)
        ns = call_ns if self.needs_ns.get(fn) else None
        if self.needs_ns.get(fn) and not ns:
            ns = fn
        key = (section, fn, ns)
        if key in self.spec:
            return self.spec[key]
        if key in self.in_progress:
            raise NotImplementedError(f"{self.key}: recursive user function {fn}")
        self.in_progress.add(key)
        _, name, params, locals_, instances, body = self.fundefs[fn]
        cn = f"fn_{section}_{re.sub(r'[^A-Za-z0-9_]', '_', name)}" + (f"__ns_{re.sub(r'[^A-Za-z0-9_]', '_', ns)}" if ns else "")
        local_map = {l: f"__fnlocal__{section}__{name}__{l}" for l in locals_}
        em = Emitter(self, section, local_map, set(params), instance_map={i: f"{ns}.{i}" for i in instances} if ns else {}, namespace=ns)
        val = em.ex(body)
        args = "".join(f", R p_{re.sub(r'[^A-Za-z0-9_]', '_', p)}" for p in params)
        code = [f"template <typename R> __device__ ZA_JSFX_FN R {cn}(Rt<R>& c, R* v{args})", "{"] + em.lines + [f"    return {val};", "}", ""]
        self.spec_code += code
        self.fn_trace[cn] = em.trace
        self.in_progress.discard(key)
        self.spec[key] = cn
        return cn


def lit(v: float) -> str:
    if v != v or v in (float("inf"), float("-inf")):
        raise ValueError("non-finite literal")
    return f"R({v!r})"


class Emitter:
    """Expression tree -> straight-line C++ statements with explicit temporaries (evaluation order = the
    reference's: left operand, right operand; right-hand side before the assignment target)."""

    def __init__(self, unit: Unit, section: str, local_map: dict, params: set, instance_map: dict = None, namespace: str = None):
        self.u, self.section, self.local_map, self.params = unit, section, local_map, params
        self.instance_map, self.namespace = instance_map or {}, namespace
        self.lines, self.n, self.ind = [], 0, 1
        # what the stage planner reads (plan_stages): ("acc", key, read, write) per variable / heap / PRNG access in evaluation
        # order, ("if",) ("else",) ("endif",) ("loop",) ("endloop",) around conditional code, ("call", fn) for user functions
        self.trace, self.stmt_lines = [], []

    def run(self, prog):
        for i, n in enumerate(prog):
            self.stmt_lines.append(len(self.lines))
            self.trace.append(("stmt", i))
            self.ex(n)
        return self.lines

    def acc(self, key, read, write):
        self.trace.append(("acc", key, read, write))

    def w(self, s):
        self.lines.append("    " * self.ind + s)

    def tmp(self, expr=None):
        self.n += 1
        t = f"t{self.n}"
        self.w(f"R {t}{'' if expr is None else ' = ' + expr};")
        return t

    # ---- variables --------------------------------------------------------------------------------------
    def ref(self, name: str, write: bool = False, read: bool = True) -> str:
        r = self._ref(name, write, read)
        if not r.startswith("p_"):
            self.acc(r.split(" /*")[0], read, write)
        return r

    def _ref(self, name: str, write: bool = False, read: bool = True) -> str:
        if name in self.params:
            return "p_" + re.sub(r"[^A-Za-z0-9_]", "_", name)
        if name in self.local_map:
            name = self.local_map[name]
        elif name in self.instance_map:                     # instance(s1) called as ns.f() -> the global pseudo-object member ns.s1
            name = self.instance_map[name]
        elif name == "this":
            name = self.namespace or name
        elif name.startswith("this."):
            suffix = name[5:]
            name = (f"{self.namespace}.{suffix}" if suffix else self.namespace) if self.namespace else (suffix or name)
        m = re.fullmatch(r"spl(\d+)", name)
        if m:
            i = int(m.group(1))
            if i >= 64:
                raise NotImplementedError("spl index >= 64")
            self.u.used_spl.add(i)
            if write:
                self.u.spl_writes.add(i)
            if read:
                self.u.spl_reads.add(i)
            return f"c.spl[{i}]"
        m = re.fullmatch(r"slider(\d+)", name)
        if m and 1 <= int(m.group(1)) <= 64:
            self.u.used_sliders.add(int(m.group(1)) - 1)
            return f"c.sliders[{int(m.group(1)) - 1}]"
        if name in STATE_BUILTINS:
            return f"c.{STATE_BUILTINS[name]}"
        if name.startswith("gmem") or name == "mem":
            raise NotImplementedError(f"{self.u.key}: {name}")
        return f"v[{self.u.slot(name)}] /*{name}*/"

    # ---- expressions ------------------------------------------------------------------------------------
    def ex(self, n) -> str:
        k = n[0]
        if k == "num":
            return lit(n[1])
        if k == "str":
            return "R(0)"
        if k == "var":
            name = n[1]
            if name == "$pi":
                return "R(3.141592653589793)"
            if name == "$e":
                return "R(2.718281828459045)"
            if name == "$phi":
                return "R(1.618033988749895)"
            if name.startswith("$x"):
                return lit(float(int(name[2:], 16)))
            if name == "mem":                      # mem[i]: the heap itself is base 0 (:4273-4275)
                return "R(0)"
            return self.tmp(self.ref(name))
        if k == "index":
            b = self.ex(n[1]); i = self.ex(n[2])
            self.acc("__mem", True, False)
            return self.tmp(f"c.ldi(c.addr1({i}))" if b == "R(0)" else f"c.ld({b}, {i})")
        if k == "un":
            a = self.ex(n[2])
            if n[1] == "+":
                return a
            if n[1] == "-":
                return self.tmp(f"sub_(R(0), {a})")
            return self.tmp(f"({a} == R(0)) ? R(1) : R(0)")
        if k == "tern" or k == "if":
            c = self.ex(n[1])
            t = self.tmp()
            self.w(f"if (truthy_({c})) {{"); self.ind += 1
            self.trace.append(("if",))
            a = self.ex(n[2]); self.w(f"{t} = {a};"); self.ind -= 1
            self.w("} else {"); self.ind += 1
            self.trace.append(("else",))
            b = self.ex(n[3]) if n[3] is not None else "R(0)"
            self.w(f"{t} = {b};"); self.ind -= 1
            self.w("}")
            self.trace.append(("endif",))
            return t
        if k == "seq":
            v = "R(0)"
            for it in n[1]:
                v = self.ex(it)
            return v
        if k == "loop":
            cnt = self.ex(n[1])
            t = self.tmp("R(0)")
            self.n += 1
            i = f"i{self.n}"
            self.w(f"for (long long {i} = 0, n{i} = to_count({cnt}); {i} < n{i}; ++{i}) {{"); self.ind += 1
            self.trace.append(("loop",))
            v = self.ex(n[2]); self.w(f"{t} = {v};"); self.ind -= 1
            self.w("}")
            self.trace.append(("endloop",))
            return t
        if k == "while":
            self.w("for (;;) {"); self.ind += 1
            self.trace.append(("loop",))
            c = self.ex(n[1]); self.w(f"if (!truthy_({c})) break;")
            self.ex(n[2]); self.ind -= 1
            self.w("}")
            self.trace.append(("endloop",))
            return "R(0)"
        if k == "bin":
            op = n[1]
            if op in ("&&", "||"):
                l = self.ex(n[2])
                t = self.tmp("R(1)" if op == "||" else "R(0)")
                self.w(f"if ({'!' if op == '||' else ''}truthy_({l})) {{"); self.ind += 1
                self.trace.append(("if",))
                r = self.ex(n[3]); self.w(f"{t} = truthy_({r}) ? R(1) : R(0);"); self.ind -= 1
                self.w("}")
                self.trace += [("else",), ("endif",)]
                return t
            l = self.ex(n[2]); r = self.ex(n[3])
            return self.tmp(self.binop(op, l, r))
        if k == "assign":
            op, tgt = n[1], n[2]
            rhs = self.ex(n[3])
            if tgt[0] == "var":
                lv = self.ref(tgt[1], write=True, read=(op != "="))
                if op == "=":
                    self.w(f"{lv} = {rhs};")
                    return rhs
                out = self.tmp(self.binop(op[:-1] if op != "~=" else "~", lv, rhs))
                self.w(f"{lv} = {out};")
                return out
            if tgt[0] == "index":
                b = self.ex(tgt[1]); i = self.ex(tgt[2])
                self.n += 1
                a = f"a{self.n}"
                self.acc("__mem", True, True)
                self.w(f"const auto {a} = c.addr1({i});" if b == "R(0)" else f"const auto {a} = c.addr({b}, {i});")
                if op == "=":
                    self.w(f"c.st({a}, {rhs});")
                    return rhs
                cur = self.tmp(f"c.ldi({a})")
                out = self.tmp(self.binop(op[:-1] if op != "~=" else "~", cur, rhs))
                self.w(f"c.st({a}, {out});")
                return out
            # slider(i) / spl(i)
            self.acc("__dyn", True, True)
            idx = self.ex(tgt[2][0])
            arr = "sliders" if tgt[1] == "slider" else "spl"
            base = "to_i32(%s) - 1" % idx if tgt[1] == "slider" else "to_i32(%s)" % idx
            self.n += 1
            a = f"a{self.n}"
            self.w(f"const int {a} = {base};")
            if op == "=":
                self.w(f"if ({a} >= 0 && {a} < 64) c.{arr}[{a}] = {rhs};")
                return rhs
            cur = self.tmp(f"({a} >= 0 && {a} < 64) ? c.{arr}[{a}] : R(0)")
            out = self.tmp(self.binop(op[:-1] if op != "~=" else "~", cur, rhs))
            self.w(f"if ({a} >= 0 && {a} < 64) c.{arr}[{a}] = {out};")
            return out
        if k == "call":
            return self.call(n[1], n[2])
        if k == "fundef":
            raise NotImplementedError("nested function definition")
        raise NotImplementedError(k)

    def binop(self, op, l, r):
        if op == "+":
            return f"add_({l}, {r})"
        if op == "-":
            return f"sub_({l}, {r})"
        if op == "*":
            return f"mul_({l}, {r})"
        if op == "/":
            return f"div_({l}, {r})"
        if op == "^":
            return f"pow_({l}, {r})"
        if op == "!=":                      # ordered, like every compare of the reference (fcmp one, dsp_jsfx_aot.py:4349-4352)
            return f"ne_({l}, {r}) ? R(1) : R(0)"
        if op in ("<", "<=", ">", ">=", "=="):
            return f"({l} {op} {r}) ? R(1) : R(0)"
        if op == "|":
            return f"R(to_i32({l}) | to_i32({r}))"
        if op == "&":
            return f"R(to_i32({l}) & to_i32({r}))"
        if op == "~":
            return f"R(to_i32({l}) ^ to_i32({r}))"
        if op == "<<":
            return f"R(shl_i32(to_i32({l}), to_i32({r})))"
        if op == ">>":
            return f"R(to_i32({l}) >> (to_i32({r}) & 31))"
        if op == "%":
            return f"R(rem_i32(to_i32({l}), to_i32({r})))"
        raise NotImplementedError(op)

    def call(self, fn, args):
        base, call_ns = self.u.resolve_call(fn, self.namespace)
        if base is not None:
            av = [self.ex(a) for a in args]
            want = len(self.u.fundefs[base][2])
            av = (av + ["R(0)"] * want)[:want]
            cn = self.u.specialise(self.section, base, call_ns)
            self.trace.append(("call", cn))
            return self.tmp(f"{cn}<R>(c, v{''.join(', ' + a for a in av)})")
        if fn in ("sliderchange", "slider_automate") and 1 <= len(args) <= 2:
            # a direct sliderN argument is its bit (_emit_slider_mask_arg, :4133-4144); both builtins OR the mask into the pending
            # masks (the gesture begin / end hint of slider_automate only concerns the host UI)
            m = re.fullmatch(r"slider([1-9][0-9]?)", args[0][1]) if args[0][0] == "var" else None
            mask = lit(float(1 << (int(m.group(1)) - 1))) if m and 1 <= int(m.group(1)) <= 64 else self.ex(args[0])
            if len(args) == 2:
                self.ex(args[1])
            self.acc("__pending", True, True)
            return self.tmp(f"c.sliderchange({mask})")
        if fn in FFT_CALLS and len(args) == 2:             # runtime hooks (:5498-5564) -> per-stream device FFT on the heap (jsfx_rt.cuh)
            av = [self.ex(a) for a in args]
            self.u.uses_fft = True
            self.acc("__mem", True, True)
            return self.tmp(FFT_CALLS[fn].format(*av))
        if fn == "convolve_c" and len(args) == 3:
            av = [self.ex(a) for a in args]
            self.u.uses_fft = True
            self.acc("__mem", True, True)
            return self.tmp(f"c.convolve_c({av[0]}, {av[1]}, {av[2]})")
        if fn.startswith("gfx_") or fn in NOOP_CALLS or fn in ("slider_show", "freembuf"):
            for a in args:
                self.ex(a)
            return "R(0)"
        if fn in UNSUPPORTED or fn.startswith(("gmem", "msg_", "sample_", "midi", "file_")) or "." in fn:
            raise NotImplementedError(f"{self.u.key}: builtin {fn}")
        av = [self.ex(a) for a in args]
        if fn in ("min", "max") and len(av) == 2:
            return self.tmp(f"({av[0]} {'<' if fn == 'min' else '>'} {av[1]}) ? {av[0]} : {av[1]}")
        if fn == "sqr" and len(av) == 1:
            return self.tmp(f"mul_({av[0]}, {av[0]})")
        if fn == "sign" and len(av) == 1:
            return self.tmp(f"({av[0]} > R(0)) ? R(1) : (({av[0]} < R(0)) ? R(-1) : R(0))")
        if fn in MATH1 and len(av) == 1:
            return self.tmp(f"m_{fn}({av[0]})")
        if fn in MATH2 and len(av) == 2:
            return self.tmp(f"m_{fn}({av[0]}, {av[1]})")
        if fn == "invsqrt" and len(av) == 1:
            return self.tmp(f"m_invsqrt({av[0]})")
        if fn == "rand" and len(av) <= 1:
            self.acc("__mt", True, True)
            return self.tmp(f"c.rand({av[0] if av else 'R(1)'}, {1 if av else 0})")
        if fn in ("memset", "memcpy") and len(av) == 3:
            self.acc("__mem", True, True)
        if fn in ("spl", "slider") and len(av) == 1:
            self.acc("__dyn", True, True)
        if fn == "memset" and len(av) == 3:
            self.w(f"c.memset({av[0]}, {av[1]}, {av[2]});")
            return av[0]
        if fn == "memcpy" and len(av) == 3:
            self.w(f"c.memcpy({av[0]}, {av[1]}, {av[2]});")
            return "R(0)"
        if fn == "spl" and len(av) == 1:
            return self.tmp(f"c.spl_dyn({av[0]})")
        if fn == "slider" and len(av) == 1:
            return self.tmp(f"c.slider_dyn({av[0]})")
        raise NotImplementedError(f"{self.u.key}: call {fn}/{len(av)}")


# ------------------------------------------------------------------------------------------------------
# stage planner: @sample cut into K consecutive groups of top-level statements, one warp per group (jsfx_rt.cuh,
# jsfx_staged_kernel).  Stage k works on frame batch i - k while stage k + 1 works on batch i - k - 1, so a value may only
# flow FORWARD within a frame; whatever a frame carries over to the next one must stay inside one stage.
# ------------------------------------------------------------------------------------------------------
STAGE_CUTS = {
    # leaf -> ([variables whose first top-level assignment opens a new stage], {special keys allowed in several stages})
    # SOMA: drive + look-ahead ring | detector | gain computer | add-back filters + presence | body saturator, motion noise,
    # sum | ceiling guards, meters, history.  The heap is touched by the first stage (look-ahead rings) and the last (history
    # arrays): disjoint regions.
    "SOMA": (["absL", "depth_now", "add_lp120L", "bodyDrive", "safety_gain"], ("__mem",)),
    # RTT: band powers | envelope followers | voice / other excess | permissions | guard + references | tail seed and age |
    # threshold + target | gain smoothing + output.  `prev_ref_off = ref_off` (last statement) moves up behind its only reader.
    "RTT": (["Ey", "v_exc", "Eo_db", "guard_tgt", "tail_seed_ratio_db", "age_w", "tgt_db_sm"], (), ("prev_ref_off",)),
    # ATTACK reuses its temporaries (s, v, oldvS) in every band block, so its cuts are statement numbers: band split | slopes of
    # the low and mid band | slopes of the high band + magnitudes | reference followers | ratios | envelopes | gains + output
    "ATTACK": ([20, 52, 80, 92, 104, 105], ()),
    # SaliencePush and ADS (no heap, no rand): cuts chosen by `tools/stage_plan.py <key> <file> --auto K`, which balances the
    # emitted lines per stage over the legal cut points
    "SaliencePush": ([20, 35, 45, 58, 69], ()),
    "ADS": ([21, 38, 56, 71, 94, 112, 130], ()),
}


class StagePlan:
    pass


def plan_stages(u: "Unit", cut_vars, n_ch: int, n_out: int, allow=(), moves=()):
    prog = u.progs["sample"]
    # per-statement pieces of the access trace
    segs, cur = [], None
    for ev in u.trace["sample"]:
        if ev[0] == "stmt":
            cur = []; segs.append(cur)
        else:
            cur.append(ev)

    def flat(trace, out):
        for ev in trace:
            if ev[0] == "acc":
                out.append(ev)
            elif ev[0] == "call":
                flat(u.fn_trace[ev[1]], out)
        return out
    # moves: the LAST top-level `x = ...` is hoisted to just behind the last earlier statement that touches x (a trailing
    # `prev = cur` would otherwise tie the first reader of prev and the end of @sample into one stage).  Legal when nothing
    # in between writes what the moved statement reads.
    order = list(range(len(prog)))
    for name in moves:
        src = [i for i, n in enumerate(prog) if n[0] == "assign" and n[1] == "=" and n[2] == ("var", name)][-1]
        key = f"v[{u.vars[name]}]"
        mine = flat(segs[src], [])
        if any(k.startswith("__") for _, k, _, _ in mine):
            raise ValueError(f"{u.key}: cannot move the assignment to {name}")
        reads = {k for _, k, r, _ in mine if r}
        pos = order.index(src)
        dst = pos
        while dst > 0 and not any(k == key for _, k, _, _ in flat(segs[order[dst - 1]], [])):
            if any(w and k in reads for _, k, _, w in flat(segs[order[dst - 1]], [])):
                raise ValueError(f"{u.key}: moving the assignment to {name} would cross a write of what it reads")
            dst -= 1
        order.insert(dst, order.pop(pos))
    cuts = []
    for name in cut_vars:
        if isinstance(name, int) or name.isdigit():        # a statement number (scripts that reuse their temporaries)
            cuts.append(order.index(int(name)))
            continue
        hit = [j for j, i in enumerate(order) if prog[i][0] == "assign" and prog[i][2] == ("var", name)]
        if not hit:
            raise ValueError(f"{u.key}: no top-level assignment to {name}")
        cuts.append(hit[0])
    if cuts != sorted(set(cuts)) or (cuts and cuts[0] == 0):
        raise ValueError(f"{u.key}: stage cuts out of order: {cuts}")
    K = len(cuts) + 1
    special = ("__mem", "__mt", "__pending")
    lw, writers, D = {}, {}, set()
    exposed, imports, special_stages = set(), {}, {k: set() for k in special}
    for ch in range(n_ch):                                 # the frame prologue of stage 0 sets every channel
        lw[f"c.spl[{ch}]"] = 0; D.add(f"c.spl[{ch}]")
    state = {"stage": 0, "cond": 0, "D": D}
    stack = []

    def access(key, r, w):
        st = state["stage"]
        if key == "__dyn":
            raise ValueError(f"{u.key}: spl(i) / slider(i) with a computed index cannot be staged")
        if key in special:
            special_stages[key].add(st)
            return
        if r or (w and state["cond"] > 0):
            if key in lw:
                if lw[key] != st:
                    imports[(key, st)] = lw[key]
            elif key not in state["D"]:
                exposed.add((st, key))
        if w:
            lw[key] = st
            writers.setdefault(key, set()).add(st)
            if state["cond"] == 0:
                state["D"].add(key)

    def walk(trace, top):
        for ev in trace:
            k = ev[0]
            if k == "acc":
                access(ev[1], ev[2], ev[3])
            elif k == "if":
                stack.append(["if", set(state["D"]), None]); state["cond"] += 1
            elif k == "else":
                stack[-1][2] = state["D"]; state["D"] = set(stack[-1][1])
            elif k == "endif":
                fr = stack.pop(); state["D"] = fr[2] & state["D"]; state["cond"] -= 1
            elif k == "loop":
                stack.append(["loop", set(state["D"])]); state["cond"] += 1
            elif k == "endloop":
                fr = stack.pop(); state["D"] = fr[1]; state["cond"] -= 1
            elif k == "call":
                walk(u.fn_trace[ev[1]], False)

    for j, i in enumerate(order):
        state["stage"] = sum(1 for c in cuts if j >= c)
        walk(segs[i], True)
    state["stage"] = K - 1
    for ch in range(n_out):                                # the frame epilogue of the last stage stores the outputs
        access(f"c.spl[{ch}]", True, False)
    own = {k: max(v) for k, v in writers.items()}
    bad = sorted(f"{k} (read at the frame start in stage {st}, written last in stage {own[k]})" for st, k in exposed if k in own and own[k] != st)
    if bad:
        raise ValueError(f"{u.key}: cut carries a value backwards: " + "; ".join(bad))
    for k, st in special_stages.items():
        if len(st) > 1 and k not in allow:
            raise ValueError(f"{u.key}: {k} is used by stages {sorted(st)}")
    for k in own:
        if k.startswith("c.sliders["):
            raise ValueError(f"{u.key}: @sample writes {k}")

    # @slider / @block run in every stage warp on that warp's copy of the variables: they may not read what @sample computes,
    # and must be repeatable (no heap stores, no rand)
    plan_block_writes = set()
    for sec in ("slider", "block"):
        for _, key, r, w in flat(u.trace[sec], []):
            if w:
                plan_block_writes.add(key)
            if key in own and r and key not in allow:
                raise ValueError(f"{u.key}: @{sec} reads {key}, which @sample writes")
            if ((key == "__mem" and w) or key in ("__mt", "__dyn")) and f"{sec}:{key}" not in allow:
                raise ValueError(f"{u.key}: @{sec} uses {key}")

    plan = StagePlan()
    plan.K, plan.cuts, plan.own, plan.order = K, cuts, own, order
    plan.block_writes = plan_block_writes
    mails = {}                                             # (key, from) -> furthest consumer
    for (key, to), frm in imports.items():
        mails[(key, frm)] = max(mails.get((key, frm), 0), to)
    plan.mail, off = {}, 0                                 # (key, from) -> (offset in slots, ring depth)
    for (key, frm), to in sorted(mails.items(), key=lambda kv: (kv[0][1], kv[0][0])):
        depth = to - frm + 1
        plan.mail[(key, frm)] = (off, depth); off += depth
    plan.slots = off
    plan.imports = [[(key, *plan.mail[(key, frm)]) for (key, to), frm in sorted(imports.items()) if to == k] for k in range(K)]
    plan.exports = [[(key, *od) for (key, frm), od in sorted(plan.mail.items()) if frm == k] for k in range(K)]
    plan.special = {k: sorted(v) for k, v in special_stages.items()}
    return plan


def emit_stages(u: "Unit", plan: StagePlan):
    out = []
    body, sl = u.body["sample"], u.stmt_lines["sample"] + [len(u.body["sample"])]
    bounds = [0] + plan.cuts + [len(u.progs["sample"])]
    moved = [i for j, i in enumerate(plan.order) if i != j and (j == 0 or plan.order[j - 1] != i - 1) and i > j]
    out.append(f"// @sample as {plan.K} pipeline stages (one warp each, jsfx_staged_kernel): stage k runs statements {bounds} [k] .. [k+1] of")
    out.append("// @sample; values that cross a cut travel through per-value rings in shared memory (mb: [slot][F frames][32 streams]; ring depth = stages crossed + 1)."
               + (f"  Statements hoisted: {moved}." if moved else ""))
    cell = lambda off, depth: f"mb[({off} + bt.r{depth}) * (F * 32) + fl]"       # bt.rD = batch number mod D
    for k in range(plan.K):
        out.append(f"template <typename R, int F> __device__ __forceinline__ void stage{k}(Rt<R>& c, R* v, R* mb, const BatchMod& bt, int fl)")
        out.append("{")
        for key, off, depth in plan.imports[k]:
            out.append(f"    {key} = {cell(off, depth)};")
        for i in plan.order[bounds[k]:bounds[k + 1]]:
            out += body[sl[i]:sl[i + 1]]
        for key, off, depth in plan.exports[k]:
            out.append(f"    {cell(off, depth)} = {key};")
        out.append("}")
        out.append("")
    out.append("struct StagedImpl {")
    out.append(f"    static constexpr int kStages = {plan.K}, kMailSlots = {max(plan.slots, 1)};")
    own = {key: 0 for key in plan.block_writes if key.startswith("v[")}
    own.update({key: st for key, st in plan.own.items() if key.startswith("v[")})
    for i0, nm in u.alias.items():                          # slider aliases are pushed per block
        own.setdefault(f"v[{u.vars[nm]}]", 0)
    sel = "".join(f"k == {int(key[2:-1])} ? {st} : " for key, st in sorted(own.items(), key=lambda kv: int(kv[0][2:-1])))
    out.append("    // the stage whose copy of v[k] is the state at the end of a call (-1: nothing on the process path writes v[k])")
    out.append(f"    static __host__ __device__ constexpr int owner(int k) {{ return {sel}-1; }}")
    out.append("    template <typename R, int F, int K> static __device__ __forceinline__ void run(Rt<R>& c, R* v, R* mb, const BatchMod& bt, int fl)")
    out.append("    {")
    for k in range(plan.K):
        out.append(f"        if constexpr (K == {k}) stage{k}<R, F>(c, v, mb, bt, fl);")
    out.append("    }")
    out.append("};")
    out.append("#undef ZA_JSFX_STAGED_IMPL")
    out.append("#define ZA_JSFX_STAGED_IMPL StagedImpl")
    out.append("")
    return out


# ------------------------------------------------------------------------------------------------------
# translation unit
# ------------------------------------------------------------------------------------------------------
def c_string(s: str) -> str:
    return '"' + s.replace("\\", "\\\\").replace('"', '\\"') + '\\n"'


def generate(key: str, jsfx_path: str, mem_slots: int, source_label: str, fft_max: int = 32768) -> str:
    """fft_max: the largest transform size the script uses (sizes the per-stream FFT scratch: 2 x fft_max slots)."""
    text = Path(jsfx_path).read_text(errors="replace")
    u = Unit(key, text)
    ident = re.sub(r"[^A-Za-z0-9_]", "_", key)
    names = sorted(u.vars, key=lambda k: u.vars[k])
    alias_slots = [-1] * 64
    for i0, nm in u.alias.items():
        alias_slots[i0] = u.vars[nm]
    nch = max(u.n_in, u.n_out, (max(u.used_spl) + 1) if u.used_spl else 0, 1)
    out = []
    out.append(f"// g_{ident}.cu -- GENERATED by za_gpu/jsfxc.py from {source_label}; do not edit.")
    out.append("// One stream per thread: the script's sections and user functions as device code (see jsfxc.py for the")
    out.append("// semantics restated from the reference's AOT compiler) + the descriptor p_generic.cu registers.")
    out.append('#include "../jsfx_rt.cuh"')
    out.append("")
    out.append(f"namespace za {{ namespace jsfx {{ namespace g_{ident} {{")
    out.append("")
    out.append(f"constexpr int kVars = {max(len(names), 1)}, kIn = {u.n_in}, kOut = {u.n_out}, kCh = {nch};")
    out.append(f"constexpr long long kFftScratch = {2 * fft_max if u.uses_fft else 0};     // slots of per-stream FFT scratch behind the heap")
    out.append("")
    out += u.spec_code
    for s in SECTIONS:
        out.append(f"template <typename R> __device__ __forceinline__ void sec_{s}(Rt<R>& c, R* v)")
        out.append("{")
        out += u.body[s]
        out.append("}")
        out.append("")
    out.append("static const char* const kVarNames[] = {" + ", ".join('"%s"' % n.replace('"', "") for n in names) + (", " if names else "") + "nullptr};")
    out.append("static const int kAlias[64] = {" + ", ".join(map(str, alias_slots)) + "};")
    sel = "".join(f"i0 == {i} ? {a} : " for i, a in enumerate(alias_slots) if a >= 0)
    out.append(f"__host__ __device__ constexpr int alias_slot(int i0) {{ return {sel}-1; }}")
    decl = [int(m.group(1)) - 1 for m in (re.match(r"\s*slider(\d+)\s*:", l) for l in u.slider_lines) if m and 1 <= int(m.group(1)) <= 64]
    out.append(f"constexpr int kDecl = {len(decl)};     // declared sliders, in declaration order (index0)")
    out.append("static const int kDeclIndex0[] = {" + ", ".join(map(str, decl or [0])) + "};")
    dsel = "".join(f"j == {j} ? {i0} : " for j, i0 in enumerate(decl))
    out.append(f"__host__ __device__ constexpr int decl_slot(int j) {{ return {dsel}0; }}")
    out.append("static const char kSliderText[] =")
    for l in u.slider_lines:
        out.append("    " + c_string(l.strip()))
    if not u.slider_lines:
        out.append('    ""')
    out.append("    ;")
    out.append("")
    if key in STAGE_CUTS:
        cut_vars, allow, *moves = STAGE_CUTS[key]
        out += emit_stages(u, plan_stages(u, cut_vars, nch, u.n_out, allow, moves[0] if moves else ()))
    out.append(f"ZA_JSFX_DEFINE_PLUGIN({ident}, \"{key}\", {mem_slots}, {1 if u.body['block'] else 0}, {1 if u.body['slider'] else 0})")
    out.append("")
    out.append("}}}  // namespace")
    out.append("")
    return "\n".join(out)


if __name__ == "__main__":
    key, path, mem = sys.argv[1], sys.argv[2], int(sys.argv[3])
    sys.stdout.write(generate(key, path, mem, sys.argv[4] if len(sys.argv) > 4 else path))
