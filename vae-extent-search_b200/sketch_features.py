"""Per-statement feature rows of a FIXED sketch as a function of its loop extents.

The reference gets the 164 features of every BufferStore by lowering each candidate with TVM on the CPU
(GetPerStoreFeaturesWorkerFunc, src/auto_scheduler/feature.cc:1296-1368) and walking the TIR with
PerStoreFeatureExtractor (:609-1048).  Within one sketch the candidates of the extent search differ only in their
tile extents: the loop structure, the buffer accesses and the operation counts of every statement are the same.  A
SketchTemplate states that structure once -- loops whose extents are products / quotients of entries of the extent
vector, buffer accesses as index expressions over the loop variables, operation counts per statement -- and
`compile()` flattens it into an int32 program that the CUDA kernel `vaex_sketch_features` (csrc/sketch_features.cu)
evaluates for every candidate: the extractor's walk (touched regions per loop level, reuse type / distance, strides,
the arithmetic-intensity curve, allocation and outer-scope features) with the candidate's extents substituted, one
164-d row per stored buffer written straight into the row matrix the SegmentSum-MLP scores.  Extents go in (4 bytes
each); no feature row crosses PCIe.

A template is data, not code: `as_dict()` is the form the test oracle (oracle/perstore_features.py) walks with the
reference's algorithm; parity tests compare the two on random extents.  What a template cannot certify is that it equals
the loop nest TVM's lowering would emit for the sketch -- TVM is not built here.
"""
import contextlib

import numpy as np
import torch

from . import _lib

ROW = 164
KINDS = {"for": 0, "vectorize": 1, "unroll": 2, "parallel": 3}
THREADS = ["blockIdx.x", "blockIdx.y", "blockIdx.z", "threadIdx.x", "threadIdx.y", "threadIdx.z", "vthread"]
OPS = ["float_mad", "float_addsub", "float_mul", "float_divmod", "float_cmp", "float_math_func", "float_other_func",
       "int_mad", "int_addsub", "int_mul", "int_divmod", "int_cmp", "int_math_func", "int_other_func", "bool_op",
       "select_op"]
K_READ, K_WRITE, K_READWRITE, K_UNKNOWN = 0, 1, 2, 3
# device program limits (csrc/sketch_features.cu)
MAX_SLOTS, MAX_SCALARS, MAX_LOOPS, MAX_BUFS, MAX_DEPTH, MAX_DIMS, MAX_TOKENS, SCALAR_SLOTS, MAX_ACC, MAX_EXPR, MAX_PROG = \
    64, 48, 32, 6, 24, 4, 32, 12, 5, 40, 4096
OP_FOR, OP_FOR_END, OP_THREAD, OP_THREAD_END, OP_PRAGMA, OP_PRAGMA_END, OP_REALIZE_END, OP_STORE = range(8)
T_VAR, T_CONST, T_ADD, T_MUL, T_DIV, T_MOD = range(6)
MAGIC = 0x56465831


def V(loop):
    """the loop variable of a loop / thread node"""
    return ("v", loop)


def C(scalar):
    return ("c", scalar)


def add(a, b):
    return ("+", a, b)


def mul(a, b):
    return ("*", a, b)


def floordiv(a, b):
    return ("//", a, b)


def floormod(a, b):
    return ("%", a, b)


TERM_PLAIN, TERM_DIV, TERM_MOD, TERM_DIVMOD, TERM_CONST = range(5)
MAX_TERMS = 8


def _terms(e):
    """The canonical form most indices have -- a left-to-right sum of terms base [* c] with base = v | v // c | v % c |
    (v // c1) % c2, or a constant -- as [(kind, loop node, coef scalar or -1, c1 scalar, c2 scalar)]; None for anything else
    (the device then walks the token form).  The device evaluates this form without an expression stack."""
    def base(x):
        if x[0] == "v":
            return (TERM_PLAIN, x[1], -1, -1)
        if x[0] in ("//", "%") and x[2][0] == "c":
            if x[1][0] == "v":
                return (TERM_DIV if x[0] == "//" else TERM_MOD, x[1][1], x[2][1], -1)
            if x[0] == "%" and x[1][0] == "//" and x[1][1][0] == "v" and x[1][2][0] == "c":
                return (TERM_DIVMOD, x[1][1][1], x[1][2][1], x[2][1])
        return None

    def term(x):
        if x[0] == "c":
            return (TERM_CONST, None, x[1], -1, -1)
        if x[0] == "*":
            if x[2][0] == "c" and x[1][0] != "c":
                b, c = base(x[1]), x[2][1]
            elif x[1][0] == "c" and x[2][0] != "c":
                b, c = base(x[2]), x[1][1]
            else:
                return None
        else:
            b, c = base(x), -1
        return None if b is None else (b[0], b[1], c, b[2], b[3])

    parts = []
    while e[0] == "+":
        parts.append(e[2])
        e = e[1]
    parts.append(e)
    out = [term(x) for x in reversed(parts)]
    return None if any(t is None for t in out) else out


def _expr_key(e):
    if e[0] == "v":
        return ("v", id(e[1]))
    if e[0] == "c":
        return e
    return (e[0], _expr_key(e[1]), _expr_key(e[2]))


class SketchTemplate:
    def __init__(self, n_slots):
        if not 0 < n_slots <= MAX_SLOTS:
            raise ValueError("1..%d extent slots" % MAX_SLOTS)
        self.n_slots = n_slots
        self.scalars, self._scalar_ix = [], {}
        self.buffers = []
        self.body = []
        self._open = [self.body]

    # ---- scalars: const * prod(extents[mul]) // prod(extents[div]) + add ----
    def scalar(self, const=1, mul=(), div=(), add=0):
        key = (int(const), tuple(sorted(mul)), tuple(sorted(div)), int(add))
        if any(not 0 <= s < self.n_slots for s in key[1] + key[2]):
            raise ValueError("slot out of range")
        if len(key[1]) + len(key[2]) > SCALAR_SLOTS:
            raise ValueError("a scalar multiplies / divides at most %d slots" % SCALAR_SLOTS)
        if key not in self._scalar_ix:
            self._scalar_ix[key] = len(self.scalars)
            self.scalars.append(key)
        return self._scalar_ix[key]

    def slot(self, j):
        return self.scalar(1, (j,))

    def const(self, v):
        return self.scalar(v)

    def buffer(self, name, shape, nbytes=4):
        if len(shape) > MAX_DIMS:
            raise ValueError("at most %d dimensions" % MAX_DIMS)
        self.buffers.append({"name": name, "shape": [int(x) for x in shape], "bytes": int(nbytes)})
        return len(self.buffers) - 1

    # ---- structure ----
    @contextlib.contextmanager
    def _scope(self, node):
        node["body"] = []
        self._open[-1].append(node)
        self._open.append(node["body"])
        try:
            yield node
        finally:
            self._open.pop()

    def loop(self, name, extent, kind="for"):
        """a For loop; extent = scalar id.  Loops whose extent evaluates to 1 vanish for that candidate (ScheduleOps drops
        trivial loops), their variable reads as 0."""
        if kind not in KINDS:
            raise ValueError(kind)
        return self._scope({"kind": kind, "name": name, "extent": extent})

    @contextlib.contextmanager
    def loops(self, *specs):
        """a perfect nest, outermost first: specs = (name, extent) or (name, extent, kind); yields the loop nodes"""
        with contextlib.ExitStack() as es:
            yield [es.enter_context(self.loop(*sp)) for sp in specs]

    def thread(self, name, extent):
        """AttrStmt thread_extent (blockIdx.* / threadIdx.*) or virtual_thread ("vthread")"""
        if name not in THREADS:
            raise ValueError(name)
        return self._scope({"kind": "thread", "name": name, "extent": extent})

    def pragma_unroll(self, value):
        """AttrStmt pragma_auto_unroll_max_step; value = scalar id"""
        return self._scope({"kind": "pragma_unroll", "value": value})

    def realize(self, buffer, bounds):
        """BufferRealize of a non-output stage at its attach point; bounds = scalar ids of the realised extents"""
        return self._scope({"kind": "realize", "buffer": buffer, "bounds": list(bounds)})

    def store(self, buffer, indices, reads=(), ops=None):
        """BufferStore buffer[indices] = f(reads); reads = [(buffer, indices), ...] in expression order; ops = counts of
        MathOpCounter (feature.cc:246-339) by name"""
        ops = dict(ops or {})
        if set(ops) - set(OPS):
            raise ValueError("unknown op counters %r" % (set(ops) - set(OPS)))
        self._open[-1].append({"kind": "store", "buffer": buffer, "indices": list(indices),
                               "reads": [(b, list(ix)) for b, ix in reads], "ops": ops})

    def as_dict(self):
        return {"slots": self.n_slots, "scalars": [(c, list(m), list(d), a) for c, m, d, a in self.scalars],
                "buffers": self.buffers, "body": self.body}

    # ---- device program ----
    def stored_buffers(self):
        order = []

        def walk(nodes):
            for n in nodes:
                if n["kind"] == "store":
                    if n["buffer"] not in order:
                        order.append(n["buffer"])
                elif "body" in n:
                    walk(n["body"])
        walk(self.body)
        return order

    @property
    def n_rows(self):
        return len(self.stored_buffers())

    def compile(self):
        """int32 program for vaex_sketch_features (layout: csrc/sketch_features.cu)"""
        if len(self.scalars) > MAX_SCALARS or len(self.buffers) > MAX_BUFS:
            raise ValueError("too many scalars / buffers for the device program")
        rows = self.stored_buffers()
        loop_id = {}
        instr, stores, exprs = [], [], []

        def expr_tokens(e, out):
            k = e[0]
            if k == "v":
                out.append(T_VAR | (loop_id[id(e[1])] << 4))
            elif k == "c":
                out.append(T_CONST | (e[1] << 4))
            else:
                expr_tokens(e[1], out)
                expr_tokens(e[2], out)
                t = {"+": T_ADD, "*": T_MUL, "//": T_DIV, "%": T_MOD}[k]
                imm = 0                                   # Mul: which operand is an immediate leaf (1 = left, 2 = right)
                if k == "*":
                    imm = 1 if e[1][0] == "c" else 2 if e[2][0] == "c" else 0
                out.append(t | (imm << 4))

        expr_cache = {}

        def expr(e):
            key = repr(_expr_key(e))
            if key in expr_cache:
                return expr_cache[key]
            toks = []
            expr_tokens(e, toks)
            if len(toks) > MAX_TOKENS:
                raise ValueError("index expression longer than %d tokens" % MAX_TOKENS)
            depth = peak = 0
            for tk in toks:                               # the device keeps the evaluation stack in four registers
                depth += 1 if (tk & 15) in (T_VAR, T_CONST) else -1
                peak = max(peak, depth)
            if peak > 4:
                raise ValueError("index expression needs an evaluation stack deeper than 4 (write sums left to right)")
            masks = [sum(1 << i for i, tk in enumerate(toks) if (tk & 15) in kinds) for kinds in ((T_ADD,), (T_MUL,), (T_DIV, T_MOD))]
            terms = _terms(e)
            tw = []
            if terms is not None and len(terms) <= MAX_TERMS:
                for kind, lp, coef, c1, c2 in terms:
                    tw.extend([kind | ((loop_id[id(lp)] if lp is not None else 0) << 4), coef, c1, c2])
            off = expr_cache[key] = len(exprs)
            exprs.extend([len(toks)] + [m - (1 << 32) if m >= (1 << 31) else m for m in masks] + [len(tw) // 4 if tw else -1] + toks + tw)
            return off

        def emit_store(n):
            acc = {n["buffer"]: [K_WRITE, [n["indices"]]]}            # BufferAccessExtractor, feature.cc:342-379
            for b, ix in n["reads"]:
                a = acc.setdefault(b, [K_UNKNOWN, []])
                if a[0] == K_WRITE:
                    a[0] = K_READWRITE
                elif a[0] == K_UNKNOWN:
                    a[0] = K_READ
                if a[0] != K_READWRITE:
                    a[1].append(ix)
            if len(acc) > MAX_ACC or sum(len(t) * len(self.buffers[b]["shape"]) for b, (_, t) in acc.items()) > MAX_EXPR:
                raise ValueError("a statement touches at most %d buffers / %d index expressions" % (MAX_ACC, MAX_EXPR))
            off = len(stores)
            stores.extend([n["buffer"], len(acc)] + [int(n["ops"].get(k, 0)) for k in OPS])
            listed = []                                                    # expression offsets in the order the device lists them
            for b, (acc_type, tuples) in acc.items():
                nd = len(self.buffers[b]["shape"])
                if any(len(ix) != nd for ix in tuples) or len(tuples) > 4:
                    raise ValueError("access to %s: %d-d indices expected, at most 4 distinct reads" % (self.buffers[b]["name"], nd))
                stores.extend([b, acc_type, len(tuples), nd])
                for ix in tuples:
                    offs = [expr(e) for e in ix]
                    stores.extend(offs)
                    listed.extend(offs)
                # one read (or the write) whose indices are all in term form: its terms, innermost loop first (a loop's id is
                # its nesting depth), constants last -- the device then takes the access through all loop levels in one walk
                # with the per-dimension widths in registers.  entry = loop id + 1 (0: constant) | dim << 8 | term << 12
                walk_ = []
                if len(tuples) == 1:
                    per_dim = [_terms(e) for e in tuples[0]]
                    if all(tm is not None and len(tm) <= MAX_TERMS for tm in per_dim):
                        for d_, tm in enumerate(per_dim):
                            for ti, (kind, lp, _, _, _) in enumerate(tm):
                                walk_.append(((loop_id[id(lp)] + 1) if kind != TERM_CONST else 0, d_, ti))
                        walk_.sort(key=lambda w_: -w_[0])
                    else:
                        walk_ = None
                else:
                    walk_ = None
                if walk_ is None:
                    stores.append(-1)
                else:
                    stores.append(len(walk_))
                    stores.extend(l_ | (d_ << 8) | (ti << 12) for l_, d_, ti in walk_)
            # every BufferLoad's indices: their arithmetic counts as int ops (MathOpCounter): (position in the list above or
            # -1, expression offset) per index
            loads = [expr(e) for _, ix in n["reads"] for e in ix]
            stores.append(len(loads))
            for o in loads:
                stores.extend([listed.index(o) if o in listed else -1, o])
            return off

        def walk(nodes, depth=0):
            for n in nodes:
                k = n["kind"]
                if k in KINDS or k == "thread":
                    # a loop's id is its nesting depth: unique among the loops that enclose a statement, which is all an
                    # index expression can name (ids are bits of a 32-bit mask on the device)
                    lid = loop_id[id(n)] = depth
                    if k == "thread":
                        instr.append((OP_THREAD, lid, n["extent"], THREADS.index(n["name"])))
                        walk(n["body"], depth + 1)
                        instr.append((OP_THREAD_END, lid, 0, THREADS.index(n["name"])))
                    else:
                        instr.append((OP_FOR, lid, n["extent"], KINDS[k]))
                        walk(n["body"], depth + 1)
                        instr.append((OP_FOR_END, lid, 0, KINDS[k]))
                elif k == "pragma_unroll":
                    instr.append((OP_PRAGMA, 0, n["value"], 0))
                    walk(n["body"], depth)
                    instr.append((OP_PRAGMA_END, 0, 0, 0))
                elif k == "realize":
                    walk(n["body"], depth)
                    if len(n["bounds"]) > MAX_DIMS:
                        raise ValueError("realize bounds")
                    off = len(stores)
                    stores.extend([len(n["bounds"])] + list(n["bounds"]))
                    instr.append((OP_REALIZE_END, n["buffer"], off, 0))
                else:
                    instr.append([OP_STORE, emit_store(n), 0, n["buffer"]])

        def depth_of(nodes):
            return max([0] + [(1 if n["kind"] in KINDS or n["kind"] == "thread" else 0) + depth_of(n["body"])
                              for n in nodes if "body" in n])

        if depth_of(self.body) > MAX_DEPTH:
            raise ValueError("loops nested deeper than %d" % MAX_DEPTH)
        walk(self.body)
        seen = set()
        for ins in reversed(instr):                       # the features a buffer keeps are those of its LAST store
            if ins[0] == OP_STORE:
                if ins[3] not in seen:
                    seen.add(ins[3])
                    ins[2] = 1
                ins[3] = 0
        sc = []
        for c, m, d, a in self.scalars:
            if not (-2 ** 31 <= c < 2 ** 31 and -2 ** 31 <= a < 2 ** 31):
                raise ValueError("scalar constant out of int32 range")
            sc.extend([c, len(m), len(d), a] + list(m) + list(d) + [0] * (SCALAR_SLOTS - len(m) - len(d)))
        bf = []
        for i, b in enumerate(self.buffers):
            bf.extend([b["bytes"], len(b["shape"])] + b["shape"] + [1] * (MAX_DIMS - len(b["shape"])) + [rows.index(i) if i in rows else -1])
        head = 12
        off_sc = head
        off_bf = off_sc + len(sc)
        off_in = off_bf + len(bf)
        off_st = off_in + 4 * len(instr)
        off_ex = off_st + len(stores)
        total = off_ex + len(exprs)
        prog = [MAGIC, self.n_slots, len(self.scalars), len(self.buffers), len(loop_id), len(rows), len(instr), off_sc, off_bf,
                off_in, off_st, off_ex]
        prog[4] = MAX_DEPTH
        prog += sc + bf + [w for ins in instr for w in ins] + stores + exprs
        assert len(prog) == total
        if total > MAX_PROG:
            raise ValueError("sketch program of %d words (limit %d)" % (total, MAX_PROG))
        return np.asarray(prog, dtype=np.int32)


class SketchFeatures:
    """A compiled template resident on the device.  rows(extents) -> float32 [n * n_rows, 164] in candidate-major order
    (candidate i owns rows [i * n_rows, (i + 1) * n_rows)), i.e. a SegmentSum-MLP batch with every segment n_rows long."""

    def __init__(self, template, device="cuda"):
        self.template = template
        self.n_rows, self.n_slots = template.n_rows, template.n_slots
        self.program = torch.from_numpy(template.compile()).to(device)

    def rows(self, extents, out=None, half=False):
        """half: fp16 rows (cvt.rn.satfinite) for the row encoder's VAEX_F_FEATS_F16 input -- same scores, half the bytes"""
        e = extents
        if not e.is_cuda:
            raise _lib.VaexError("vaex_b200 needs CUDA tensors (no CPU fallback)")
        _lib.check_device(e)
        e = e.to(torch.int32).contiguous()
        if e.dim() != 2 or e.shape[1] != self.n_slots:
            raise ValueError("extents must be [n, %d]" % self.n_slots)
        n = e.shape[0]
        if out is None:
            out = torch.empty(n * self.n_rows, ROW, device=e.device, dtype=torch.float16 if half else torch.float32)
        half = out.dtype == torch.float16
        fn = _lib.load().vaex_sketch_features_f16 if half else _lib.load().vaex_sketch_features
        _lib.check(fn(_lib.ptr(self.program), self.program.numel(), _lib.ptr(e), n, _lib.ptr(out), _lib.stream_ptr()))
        return out

    def segment_sizes(self, n, device="cuda"):
        return torch.full((n,), self.n_rows, dtype=torch.int32, device=device)


# ---------------------------------------------------------------------------------------------------------------
# templates
# ---------------------------------------------------------------------------------------------------------------
def tiled(t, parts):
    """index of an axis split into loops `parts` = [(loop, inner_scalar_or_None), ...] outermost first:
    ((p0 * c0) + (p1 * c1)) + p_last  -- the sum a split chain lowers to (outer * factor + inner, distributed)"""
    e = None
    for lp, c in parts:
        term = V(lp) if c is None else mul(V(lp), C(c))
        e = term if e is None else add(e, term)
    return e


def matmul_cpu_template(N, M, K):
    """C[i, j] = sum_k A[i, k] * B[k, j] under the schedule of the reference's feature test
    (tests/python/unittest/test_auto_scheduler_feature.py:31-75): i -> (i.0, i.1), j -> (j.0, j.1), order
    (i.0, j.0, k, j.1, i.1), i.0 / j.0 parallel, k unrolled, j.1 vectorised.  Extent vector: [i.0, j.0, k, j.1, i.1]
    (walk order of the printed state).  The reduction's init statement sits in front of the k loop with its own copies
    of the inner spatial loops (te::MakeReduction)."""
    t = SketchTemplate(5)
    A, B, Cb = t.buffer("A", (N, K)), t.buffer("B", (K, M)), t.buffer("C", (N, M))
    e_i0, e_j0, e_k, e_j1, e_i1 = (t.slot(j) for j in range(5))
    with t.loop("i.0", e_i0, "parallel") as i0:
        with t.loop("j.0", e_j0, "parallel") as j0:
            with t.loop("j.1.init", e_j1, "vectorize") as j1i:
                with t.loop("i.1.init", e_i1) as i1i:
                    t.store(Cb, [tiled(t, [(i0, e_i1), (i1i, None)]), tiled(t, [(j0, e_j1), (j1i, None)])])
            with t.loop("k", e_k, "unroll") as k:
                with t.loop("j.1", e_j1, "vectorize") as j1:
                    with t.loop("i.1", e_i1) as i1:
                        ci, cj = tiled(t, [(i0, e_i1), (i1, None)]), tiled(t, [(j0, e_j1), (j1, None)])
                        t.store(Cb, [ci, cj], reads=[(Cb, [ci, cj]), (A, [ci, V(k)]), (B, [V(k), cj])],
                                ops={"float_addsub": 1, "float_mul": 1})
    return t


def lin(terms):
    """((t0 * c0) + (t1 * c1)) + ... : terms = [(expr, coefficient scalar or None)], outermost first"""
    e = None
    for x, c in terms:
        term = x if c is None else mul(x, C(c))
        e = term if e is None else add(e, term)
    return e


# slots of conv2d_nhwc_cpu_template: four tile levels of every spatial axis, two of every reduction axis, the unroll pragma
CONV2D_SLOTS = {"n": (0, 1, 2, 3), "y": (4, 5, 6, 7), "x": (8, 9, 10, 11), "f": (12, 13, 14, 15), "ry": (16, 17), "rx": (18, 19),
                "rc": (20, 21), "unroll": 22}
# the schedule printed in the reference README (:86-108): 1 x 56 x 56 x 64 -> 1 x 28 x 28 x 128, 3 x 3, stride 2
CONV2D_README_EXTENTS = [1, 1, 1, 1, 1, 1, 1, 28, 4, 7, 1, 1, 2, 8, 1, 8, 1, 3, 3, 1, 4, 16, 16]


def conv2d_nhwc_cpu_template(N=1, H=56, W=56, CI=64, CO=128, KH=3, KW=3, stride=2, pad_before=1, pad_after=0, derive_outer=False,
                             pad_at="fused", pad_vectorize=False):
    """conv2d NHWC + bias + ReLU under Ansor's CPU multi-level tiling sketch ("SSRSRS" with the consumer fused at the second
    tile level, the padding stage computed at the fused outer loop, the outer spatial tiles fused and parallelised, the
    innermost channel tile vectorised): task 1264 of the reference README -- its compute DAG (README.md:64-76) gives the
    statements (PaddedInput with its four comparisons, the Conv2dOutput update, T_add inlined into T_relu), the program it
    prints (:86-108) the loop structure; that program's extents are CONV2D_README_EXTENTS here.  Extent vector (23 slots, CONV2D_SLOTS): the
    tile factors n, y, x, f (4 levels each, outermost first), ry, rx, rc (2 levels each) and auto_unroll_max_step.  Three
    statement rows per candidate.  derive_outer: the outermost level of every axis is not read from the vector but derived
    as axis length // product of the inner levels -- the form a SplitStep has (it stores the inner lengths only,
    transform_step.cc:1015-1023) and the divisor legaliser guarantees to be exact (conv2d_slot_tables).  pad_at / pad_vectorize:
    where the padding stage is computed ("fused": under the fused outer loop, as in the README program; "ax3.0": one loop
    further in, as in the reference's ci_logs record -- Ansor samples this location per program) and whether its channel loop
    is vectorised."""
    if pad_at not in ("fused", "ax3.0"):
        raise ValueError(pad_at)
    pad = pad_before + pad_after
    OH, OW = (H + pad - KH) // stride + 1, (W + pad - KW) // stride + 1
    t = SketchTemplate(23)
    length = {0: N, 4: OH, 8: OW, 12: CO, 16: KH, 18: KW, 20: CI}         # first slot of an axis -> the axis length
    data, kernel, bias = t.buffer("placeholder", (N, H, W, CI)), t.buffer("placeholder", (KH, KW, CI, CO)), t.buffer("placeholder", (1, 1, 1, CO))
    Pad, Conv, Relu = (t.buffer("PaddedInput", (N, H + pad, W + pad, CI)), t.buffer("Conv2dOutput", (N, OH, OW, CO)),
                       t.buffer("T_relu", (N, OH, OW, CO)))
    sl = CONV2D_SLOTS
    groups = {g[0]: g for g in (sl[k] for k in ("n", "y", "x", "f", "ry", "rx", "rc"))}

    def S(*slots, const=1, add=0):
        mul_, div_ = [], []
        for j in slots:
            if derive_outer and j in length:
                const *= length[j]
                div_ += list(groups[j][1:])
            else:
                mul_.append(j)
        for j in list(mul_):                                              # an inner level both multiplied and divided cancels
            if j in div_:
                mul_.remove(j)
                div_.remove(j)
        return t.scalar(const, tuple(mul_), tuple(div_), add)

    n, y, x, f, ry, rx, rc = (sl[k] for k in ("n", "y", "x", "f", "ry", "rx", "rc"))
    zero = t.const(0)
    with t.loop("ax0.0@ax1.0@ax2.0@", S(n[0], y[0], x[0]), "parallel") as p:
        n0 = floordiv(V(p), C(S(y[0], x[0])))
        y0 = floormod(floordiv(V(p), C(S(x[0]))), C(S(y[0])))
        x0 = floormod(V(p), C(S(x[0])))
        rows, cols = S(*y[1:], const=stride, add=KH - stride), S(*x[1:], const=stride, add=KW - stride)

        def padding_stage():
          with t.realize(Pad, [S(*n[1:]), rows, cols, t.const(CI)]):
            with t.loops(("i0", S(*n[1:])), ("i1", rows), ("i2", cols),
                         ("i3", t.const(CI), "vectorize" if pad_vectorize else "for")) as (i0, i1, i2, i3):
                  ix = [lin([(n0, S(*n[1:])), (V(i0), None)]), lin([(y0, S(*y[1:], const=stride)), (V(i1), None)]),
                        lin([(x0, S(*x[1:], const=stride)), (V(i2), None)]), V(i3)]
                  # PaddedInput(i0, i1, i2, i3) = tir.if_then_else(i1 >= 1 && i1 < 57 && i2 >= 1 && i2 < 57,
                  #                                                   placeholder[i0, i1 - 1, i2 - 1, i3], 0f)   (README.md:70)
                  shift = t.const(-pad_before)
                  src = [ix[0], add(ix[1], C(shift)), add(ix[2], C(shift)), ix[3]] if pad_before else ix
                  t.store(Pad, ix, reads=[(data, src)], ops={"float_math_func": 1, "int_cmp": 4, "bool_op": 3})
        if pad_at == "fused":
            padding_stage()
        with t.loop("ax3.0", S(f[0])) as f0:
          if pad_at == "ax3.0":
            padding_stage()
          with t.loops(("ax0.1", S(n[1])), ("ax1.1", S(y[1])), ("ax2.1", S(x[1])), ("ax3.1", S(f[1]))) as (n1, y1, x1, f1):
            outer = lambda ax, v0, v1: [(v0, S(*ax[1:])), (V(v1), S(*ax[2:]))]
            fo = [(V(f0), S(*f[1:])), (V(f1), S(*f[2:]))]
            with t.realize(Conv, [S(*n[2:]), S(*y[2:]), S(*x[2:]), S(*f[2:])]):
                with t.pragma_unroll(S(sl["unroll"])):
                    with t.loops(("nn.2.init", S(n[2])), ("yy.2.init", S(y[2])), ("xx.2.init", S(x[2])), ("ff.2.init", S(f[2])),
                                 ("nn.3.init", S(n[3])), ("yy.3.init", S(y[3])), ("xx.3.init", S(x[3])),
                                 ("ff.3.init", S(f[3]), "vectorize")) as (a2, b2, c2, d2, a3, b3, c3, d3):
                        t.store(Conv, [lin(outer(n, n0, n1) + [(V(a2), S(n[3])), (V(a3), None)]),
                                       lin(outer(y, y0, y1) + [(V(b2), S(y[3])), (V(b3), None)]),
                                       lin(outer(x, x0, x1) + [(V(c2), S(x[3])), (V(c3), None)]),
                                       lin(fo + [(V(d2), S(f[3])), (V(d3), None)])])
                    with t.loops(("ry.0", S(ry[0])), ("rx.0", S(rx[0])), ("rc.0", S(rc[0])), ("nn.2", S(n[2])), ("yy.2", S(y[2])),
                                 ("xx.2", S(x[2])), ("ff.2", S(f[2])), ("ry.1", S(ry[1])), ("rx.1", S(rx[1])), ("rc.1", S(rc[1])),
                                 ("nn.3", S(n[3])), ("yy.3", S(y[3])), ("xx.3", S(x[3])), ("ff.3", S(f[3]), "vectorize")) as \
                            (r0, s0, q0, a2, b2, c2, d2, r1, s1, q1, a3, b3, c3, d3):
                        nn = lin(outer(n, n0, n1) + [(V(a2), S(n[3])), (V(a3), None)])
                        yy = lin(outer(y, y0, y1) + [(V(b2), S(y[3])), (V(b3), None)])
                        xx = lin(outer(x, x0, x1) + [(V(c2), S(x[3])), (V(c3), None)])
                        ff = lin(fo + [(V(d2), S(f[3])), (V(d3), None)])
                        kr, ks, kq = (lin([(V(r0), S(ry[1])), (V(r1), None)]), lin([(V(s0), S(rx[1])), (V(s1), None)]),
                                      lin([(V(q0), S(rc[1])), (V(q1), None)]))
                        py = lin([(y0, S(*y[1:], const=stride)), (V(y1), S(*y[2:], const=stride)), (V(b2), S(y[3], const=stride)),
                                  (V(b3), t.const(stride)), (V(r0), S(ry[1])), (V(r1), None)])
                        px = lin([(x0, S(*x[1:], const=stride)), (V(x1), S(*x[2:], const=stride)), (V(c2), S(x[3], const=stride)),
                                  (V(c3), t.const(stride)), (V(s0), S(rx[1])), (V(s1), None)])
                        t.store(Conv, [nn, yy, xx, ff], reads=[(Conv, [nn, yy, xx, ff]), (Pad, [nn, py, px, kq]), (kernel, [kr, ks, kq, ff])],
                                ops={"float_addsub": 1, "float_mul": 1})
            with t.loops(("ax0.2", S(*n[2:])), ("ax1.2", S(*y[2:])), ("ax2.2", S(*x[2:])), ("ax3.2", S(*f[2:]), "vectorize")) as (a, b, c, d):
                ix = [lin(outer(n, n0, n1) + [(V(a), None)]), lin(outer(y, y0, y1) + [(V(b), None)]), lin(outer(x, x0, x1) + [(V(c), None)]),
                      lin(fo + [(V(d), None)])]
                # T_relu = max(Conv2dOutput[ax0, ax1, ax2, ax3] + placeholder[ax0, 0, 0, ax3], 0f)  (T_add inlined; README.md:74-75)
                t.store(Relu, ix, reads=[(Conv, ix), (bias, [ix[0], C(zero), C(zero), ix[3]])], ops={"float_addsub": 1, "float_cmp": 1})
    return t


def conv2d_slot_tables(N=1, H=56, W=56, CI=64, CO=128, KH=3, KW=3, stride=2, pad_before=1, pad_after=0, max_innermost=64):
    """(slot_extent int32[23], slot_last uint8[23]) for vaex_decode_snap / vaex_legalise_extents in divisor mode with
    conv2d_nhwc_cpu_template(derive_outer=True): the inner levels of an axis are the lengths of its SplitStep (their product
    must divide the axis length, the innermost one at most max_innermost); the outermost levels and the unroll slot are free."""
    pad = pad_before + pad_after
    OH, OW = (H + pad - KH) // stride + 1, (W + pad - KW) // stride + 1
    ext, last = np.zeros(23, np.int32), np.zeros(23, np.uint8)
    for key, E in (("n", N), ("y", OH), ("x", OW), ("f", CO), ("ry", KH), ("rx", KW), ("rc", CI)):
        g = CONV2D_SLOTS[key]
        ext[list(g[1:])] = E
        last[g[-1]] = 1
    return ext, last


class SketchScorer:
    """The north-star scoring chain for one sketch, device-resident end to end:

        z [n, L] -> VAE decoder -> snap -> divisor legalise -> extents [n, slots]          (vaex_decode_snap)
                 -> the sketch's per-statement feature rows [n * rows, 164]                 (vaex_sketch_features)
                 -> SegmentSum-MLP score per candidate [n]                                  (vaex_segsum_mlp_forward)

    The reference runs the middle step on the CPU through TVM and moves 656 bytes per statement to the cost model; here
    only the latents (or nothing, when they are produced on the device) cross PCIe.  vae: VAE_feature_head whose output
    width is the template's slot count; cost_model: SegmentSumMLPModule; mean / scale: the extent scaler; slot_extent /
    slot_last: the legaliser's tables for the template (e.g. conv2d_slot_tables())."""

    def __init__(self, vae, mean, scale, template, cost_model, slot_extent=None, slot_last=None, log1p=True, max_innermost=64,
                 fea_norm_vec=None):
        dev = next(vae.parameters()).device
        if dev.type != "cuda":
            raise _lib.VaexError("vaex_b200 needs the models on a CUDA device (no CPU fallback)")
        if vae.decoder[-1].out_features != template.n_slots:
            raise ValueError("the decoder emits %d extents, the template reads %d" % (vae.decoder[-1].out_features, template.n_slots))
        self.vae, self.cost_model, self.features = vae, cost_model, SketchFeatures(template, dev)
        self.mean = torch.as_tensor(mean, dtype=torch.float32).to(dev).contiguous()
        self.scale = torch.as_tensor(scale, dtype=torch.float32).to(dev).contiguous()
        self.slot_extent = None if slot_extent is None else torch.as_tensor(slot_extent, dtype=torch.int32).to(dev).contiguous()
        self.slot_last = None if slot_last is None else torch.as_tensor(slot_last, dtype=torch.uint8).to(dev).contiguous()
        self.mode = (_lib.SNAP_EXPM1 if log1p else _lib.SNAP_EXP) | (_lib.SNAP_DIVISOR if slot_extent is not None else 0)
        self.max_innermost, self.fea_norm_vec = max_innermost, fea_norm_vec
        self._buf = {}

    def _buffers(self, n, dev):
        b = self._buf.get(n)
        if b is None:
            R = self.features.n_rows
            b = self._buf[n] = (torch.empty(n, self.features.n_slots, dtype=torch.int32, device=dev),
                                torch.empty(n * R, ROW, device=dev, dtype=torch.float16),       # the row encoder's fp16 input
                                torch.arange(0, (n + 1) * R, R, dtype=torch.int32, device=dev),
                                torch.empty(n, device=dev))
            if len(self._buf) > 4:
                self._buf.pop(next(iter(self._buf)))
        return b

    def score(self, z):
        """-> (scores fp32 [n], extents int32 [n, slots]); buffers are reused between calls of the same n"""
        if not z.is_cuda:
            raise _lib.VaexError("vaex_b200 needs CUDA tensors (no CPU fallback)")
        _lib.check_device(z)
        z = z.detach().contiguous().float()
        n, lib, st = z.shape[0], _lib.load(), _lib.stream_ptr()
        ext, rows, offsets, scores = self._buffers(n, z.device)
        if n == 0:
            return scores, ext
        _lib.check(lib.vaex_decode_snap(self.vae.packed().h, _lib.ptr(z), n, None, _lib.ptr(ext), _lib.ptr(self.mean), _lib.ptr(self.scale),
                                        self.mode, _lib.ptr(self.slot_extent), _lib.ptr(self.slot_last), self.max_innermost, st))
        self.features.rows(ext, out=rows)
        _lib.check(lib.vaex_segsum_mlp_forward(self.cost_model.packed(self.fea_norm_vec).h, _lib.ptr(rows), _lib.ptr(offsets), n,
                                               rows.shape[0], _lib.ptr(scores), _lib.F_FEATS_F16, st))
        return scores, ext


# slots of matmul_mlt_cpu_template: four tile levels of i and j, two of k, the unroll pragma
MATMUL_SLOTS = {"i": (0, 1, 2, 3), "j": (4, 5, 6, 7), "k": (8, 9), "unroll": 10}


def matmul_mlt_cpu_template(M=512, N=512, K=512, derive_outer=False):
    """C[i, j] = sum_k A[i, k] * B[k, j] under Ansor's CPU sketch for an output stage without a consumer: multi-level tiling
    "SSRSRS" on a cache-write stage C.local computed at the second tile level of C, the two outer tile levels of C fused and
    parallelised, the innermost j tile vectorised in both stages (the CPU counterpart of the loop nest printed at
    tests/python/unittest/test_auto_scheduler_feature.py:173-195).  Extent vector (11 slots, MATMUL_SLOTS): i and j in four
    levels, k in two, auto_unroll_max_step.  Two statement rows per candidate (C.local, C)."""
    t = SketchTemplate(11)
    A, B = t.buffer("A", (M, K)), t.buffer("B", (K, N))
    Cl, Cb = t.buffer("C.local", (M, N)), t.buffer("C", (M, N))
    sl = MATMUL_SLOTS
    length = {0: M, 4: N, 8: K}
    groups = {g[0]: g for g in (sl["i"], sl["j"], sl["k"])}

    def S(*slots, const=1, add=0):
        mul_, div_ = [], []
        for s_ in slots:
            if derive_outer and s_ in length:
                const *= length[s_]
                div_ += list(groups[s_][1:])
            else:
                mul_.append(s_)
        for s_ in list(mul_):
            if s_ in div_:
                mul_.remove(s_)
                div_.remove(s_)
        return t.scalar(const, tuple(mul_), tuple(div_), add)

    i, j, k = sl["i"], sl["j"], sl["k"]
    with t.loop("i.0@j.0@", S(i[0], j[0]), "parallel") as p:
        i0, j0 = floordiv(V(p), C(S(j[0]))), floormod(V(p), C(S(j[0])))
        with t.loops(("i.1", S(i[1])), ("j.1", S(j[1]))) as (i1, j1):
            io = [(i0, S(*i[1:])), (V(i1), S(*i[2:]))]
            jo = [(j0, S(*j[1:])), (V(j1), S(*j[2:]))]
            with t.realize(Cl, [S(*i[2:]), S(*j[2:])]):
                with t.pragma_unroll(S(sl["unroll"])):
                    with t.loops(("i_c.2.init", S(i[2])), ("j_c.2.init", S(j[2])), ("i_c.3.init", S(i[3])),
                                 ("j_c.3.init", S(j[3]), "vectorize")) as (a2, b2, a3, b3):
                        t.store(Cl, [lin(io + [(V(a2), S(i[3])), (V(a3), None)]), lin(jo + [(V(b2), S(j[3])), (V(b3), None)])])
                    with t.loops(("k.0", S(k[0])), ("i_c.2", S(i[2])), ("j_c.2", S(j[2])), ("k.1", S(k[1])), ("i_c.3", S(i[3])),
                                 ("j_c.3", S(j[3]), "vectorize")) as (k0, a2, b2, k1, a3, b3):
                        ci = lin(io + [(V(a2), S(i[3])), (V(a3), None)])
                        cj = lin(jo + [(V(b2), S(j[3])), (V(b3), None)])
                        kk = lin([(V(k0), S(k[1])), (V(k1), None)])
                        t.store(Cl, [ci, cj], reads=[(Cl, [ci, cj]), (A, [ci, kk]), (B, [kk, cj])], ops={"float_addsub": 1, "float_mul": 1})
            with t.loops(("i.2", S(*i[2:])), ("j.2", S(*j[2:]), "vectorize")) as (a, b):
                ix = [lin(io + [(V(a), None)]), lin(jo + [(V(b), None)])]
                t.store(Cb, ix, reads=[(Cl, ix)])
    return t


def conv2d_template_from_record(record):
    """(template, extents) for one measure record of the conv2d NHWC + bias + ReLU workload function (the records of
    `tutorials/auto_scheduler/ci_logs/resnet-50-NHWC-B1-llvm.json`, TenSet's task 1264, ...): the shapes from the workload key
    [hash, N, H, W, CI, KH, KW, CI, CO, 1, 1, 1, CO, N, OH, OW, CO], the tile extents from the seven SplitSteps of the
    Conv2dOutput stage (their lengths are the inner levels; the outermost is extent // product), the unroll pragma, and the
    two things Ansor samples per program about the padding stage: where it is attached and whether its channel loop is
    vectorised.  Raises ValueError for a record of another sketch (different step pattern)."""
    import json
    rec = json.loads(record) if isinstance(record, (str, bytes)) else record
    key = json.loads(rec["i"][0][0])
    steps = rec["i"][1][1]
    if len(key) != 17:
        raise ValueError("not a conv2d NHWC + bias + ReLU workload key")
    N, H, W, CI, KH, KW, _, CO = key[1:9]
    OH, OW = key[14], key[15]
    stride = max(1, round(H / OH))
    total = (OH - 1) * stride + KH - H
    sp = [s_ for s_ in steps if s_[0] == "SP" and s_[1] == 3]
    if len(sp) != 7 or [len(s_[4]) for s_ in sp] != [3, 3, 3, 3, 1, 1, 1] or [s_[3] for s_ in sp] != [N, OH, OW, CO, KH, KW, CI]:
        raise ValueError("the record's Conv2dOutput stage is not tiled S S R S R S (4 x 3 lengths, 3 x 1)")
    reorder = [s_ for s_ in steps if s_[0] == "RE" and s_[1] == 3]
    if not reorder or reorder[0][2] != [0, 4, 8, 12, 1, 5, 9, 13, 16, 18, 20, 2, 6, 10, 14, 17, 19, 21, 3, 7, 11, 15]:
        raise ValueError("the record's loop order is not the template's")
    if ["CA", 3, 6, 7] not in steps or ["FU", 6, [0, 1, 2]] not in steps:
        raise ValueError("the record attaches / fuses the stages differently from the template")
    ext = []
    for s_ in sp:
        inner = [int(v) for v in s_[4]]
        ext += [int(s_[3]) // int(np.prod(inner))] + inner
    pr = [s_ for s_ in steps if s_[0] == "PR" and s_[1] == 3]
    ext.append(int(pr[0][3].split("$")[1]) if pr else 0)
    pad_ca = [s_ for s_ in steps if s_[0] == "CA" and s_[1] == 1 and s_[2] == 6]
    if not pad_ca or pad_ca[0][3] not in (2, 3):
        raise ValueError("the padding stage is attached somewhere this template does not state")
    pad_at = "ax3.0" if pad_ca[0][3] == 3 else "fused"        # iterators of the consumer before the fuse: n.0 y.0 x.0 f.0 ...
    t = conv2d_nhwc_cpu_template(N, H, W, CI, CO, KH, KW, stride, (total + 1) // 2, total // 2, pad_at=pad_at,
                                 pad_vectorize=["AN", 1, 3, 2] in steps)
    return t, ext


def conv2d_sketch_slots(record):
    """The extent.SketchSlots of the 23-slot vector of conv2d_nhwc_cpu_template for one measure record of that sketch: the
    inner tile levels are the lengths of the record's SplitSteps, slot 22 its unroll pragma, the outermost levels derived.
    With it the candidates the chain ranks highest go back out as measure-record lines (SketchSlots.records_from_extents),
    the form the reference's search loop hands to TVM for measurement."""
    import json
    from .extent import SketchSlots
    rec = json.loads(record) if isinstance(record, (str, bytes)) else record
    steps = rec["i"][1][1]
    sp = [k for k, s_ in enumerate(steps) if s_[0] == "SP" and s_[1] == 3]
    pr = [k for k, s_ in enumerate(steps) if s_[0] == "PR" and s_[1] == 3]
    if len(sp) != 7:
        raise ValueError("not a record of the conv2d sketch")
    slots = [None] * 23
    for key, k in zip(("n", "y", "x", "f", "ry", "rx", "rc"), sp):
        for level, j in enumerate(CONV2D_SLOTS[key][1:]):
            slots[j] = ("SP", k, level)
    if pr:
        slots[22] = ("PR", pr[0])
    return SketchSlots(rec, slots)
