"""TEST INFRASTRUCTURE ONLY -- CPU restatement of TVM auto_scheduler's per-store feature extractor
(src/auto_scheduler/feature.cc:60-1188) on a small loop-nest IR.

The reference obtains the 164 features of every BufferStore by lowering a schedule with TVM (ApplySteps -> InferBound ->
ScheduleOps -> Simplify, feature.cc:1296-1368) and walking the TIR.  TVM is not built in this image, so the walk is
restated here over an explicit loop nest: the *template* of a sketch (the loop structure, the buffer accesses and the
operation counts of every statement, with loop extents given as products of the entries of an extent vector -- see
vae-extent-search_b200/sketch_features.py for the format).  evaluate(template, extents) returns one 164-d row per stored
buffer, the vector GetPerStoreFeature (:1053-1188) emits for that loop nest.

Pinning: the known-answer tests of the reference for this extractor, tests/python/unittest/test_auto_scheduler_feature.py
(:31-95 cpu matmul, :97-137 fusion, :139-204 gpu matmul), whose comments spell out the lowered loop nests -- checked in
tests/test_oracle_perstore.py.  What stays UNPINNED is the step this module does not perform: that a template equals what
TVM's lowering emits for a given sketch (no TVM here to lower one).

Known order dependences of the reference (iteration over an unordered_map keyed by buffer pointers) are resolved here by
insertion order (the stored buffer first, then reads in expression order): the buffer whose region is left in
`tmp_region` for n_continuous (:881-889), ties of the buffer sort key (:1115-1120) and the order of the rows.
"""
import math

import numpy as np

F32 = np.float32
ARITH_N = 10                      # ARITH_INTENSITY_CURVE_SAMPLE_N, feature.cc:65
MAX_N_BUFS = 5                    # python/tvm/auto_scheduler/feature.py:39
CACHE_LINE = 64                   # HardwareParams default cache_line_bytes
ROW = 164
K_READ, K_WRITE, K_READWRITE, K_UNKNOWN = 0, 1, 2, 3                     # BufferAccessType, :81
R_LOOP_MULTIPLE_READ, R_SERIAL_MULTIPLE_RW, R_NO_REUSE = 0, 1, 2         # ReuseType, :93
POS_NONE, POS_MIXED = 0, 7                                               # AnnotationPosType, :68-78
OPS = ["float_mad", "float_addsub", "float_mul", "float_divmod", "float_cmp", "float_math_func", "float_other_func",
       "int_mad", "int_addsub", "int_mul", "int_divmod", "int_cmp", "int_math_func", "int_other_func", "bool_op",
       "select_op"]
THREADS = ["blockIdx.x", "blockIdx.y", "blockIdx.z", "threadIdx.x", "threadIdx.y", "threadIdx.z", "vthread"]


def scalar_values(template, extents):
    """S_k = const * prod(extents[mul]) // prod(extents[div]) + add for every scalar of the template."""
    out = []
    for const, mul, div, add in template["scalars"]:
        v = int(const)
        for s in mul:
            v *= int(extents[s])
        for s in div:
            v //= max(int(extents[s]), 1)
        out.append(v + int(add))
    return out


def index_ops(e, counts):
    """MathOpCounter (feature.cc:246-339) also walks the indices of every BufferLoad of a statement's value: the integer
    adds, multiplies and floordiv / floormod of the (simplified) index arithmetic count as int ops."""
    if e[0] in ("v", "c"):
        return
    counts[{"+": "int_addsub", "*": "int_mul", "//": "int_divmod", "%": "int_divmod"}[e[0]]] += 1
    index_ops(e[1], counts)
    index_ops(e[2], counts)


# ---------------------------------------------------------------------------------------------------------------
# expressions: ("v", loop) | ("c", scalar) | ("+", a, b) | ("*", a, b) | ("//", a, b) | ("%", a, b)
# ---------------------------------------------------------------------------------------------------------------
def var_in_expr(e, loop):
    """VarInExpr, feature.cc:164-181."""
    if e[0] == "v":
        return e[1] is loop
    if e[0] == "c":
        return False
    return var_in_expr(e[1], loop) or var_in_expr(e[2], loop)


def int_bound(e, rng, S):
    """arith::Analyzer::const_int_bound of an index under the bound loop ranges (src/arith/const_int_bound.cc): vars
    are [min, max] of their binding, immediates a point; +, * by interval arithmetic, floordiv / floormod by a positive
    constant with the analyzer's rules."""
    k = e[0]
    if k == "v":
        return rng[id(e[1])]
    if k == "c":
        return (S[e[1]], S[e[1]])
    a, b = int_bound(e[1], rng, S), int_bound(e[2], rng, S)
    if k == "+":
        return (a[0] + b[0], a[1] + b[1])
    if k == "*":
        c = [a[0] * b[0], a[0] * b[1], a[1] * b[0], a[1] * b[1]]
        return (min(c), max(c))
    assert b[0] == b[1] and b[0] > 0, "floordiv / floormod by a positive constant only"
    if k == "//":
        return (a[0] // b[0], a[1] // b[0])
    if k == "%":
        if a[0] >= 0:
            return a if a[1] < b[0] else (0, min(a[1], b[0] - 1))
        return (0, b[0] - 1)
    raise ValueError(k)


class _Coefficient:
    """CoefficientExtractor, feature.cc:382-443 (post-order visit; only Mul and Add react)."""

    def extract(self, e, loop, S):
        self.var = self.mul = self.add = False
        self.stride = 0
        self._visit(e, loop, S)
        return 1 if (self.var and not self.mul and not self.add) else self.stride

    def _visit(self, e, loop, S):
        k = e[0]
        if k == "v":
            if e[1] is loop:
                self.var = True
                self.stride = 2           # "magic default stride", :419
            return
        if k == "c":
            return
        self._visit(e[1], loop, S)
        self._visit(e[2], loop, S)
        if k == "*" and self.var and not self.add:
            if e[1][0] == "c":
                self.mul, self.stride = True, S[e[1][1]]
            elif e[2][0] == "c":
                self.mul, self.stride = True, S[e[2][1]]
        elif k == "+" and self.var and not self.mul:
            self.add, self.stride = True, 1


def compute_stride(indices, shape, loop, S):
    """ComputeStride, feature.cc:446-468."""
    best, find = None, False
    cx = _Coefficient()
    for index in indices:
        shape_stride = 1
        for i in range(len(index) - 1, -1, -1):
            coef = cx.extract(index[i], loop, S)
            if cx.var:
                find = True
                v = abs(coef) * shape_stride
                best = v if best is None else min(best, v)
                break
            shape_stride *= shape[i]
    return best if find else 0


def compute_region(indices, rng, S):
    """ComputeRegion, feature.cc:471-499."""
    if not indices:
        return []
    out = []
    for d in range(len(indices[0])):
        lo = min(int_bound(ix[d], rng, S)[0] for ix in indices)
        hi = max(int_bound(ix[d], rng, S)[1] for ix in indices)
        out.append(hi - lo + 1)
    return out


class _Loop:
    def __init__(self, name, extent, kind):
        self.name, self.extent, self.kind = name, int(extent), kind
        self.touch = {}           # buffer id -> [(acc_type, touched elements, element bytes)]  (for_touch_regions_, :1028)


def compute_reuse(buf, indices, stack):
    """ComputeReuse, feature.cc:502-603 -> (reuse_type, reuse_dis_iter, reuse_dis_bytes, reuse_ct) in float32."""
    dis_iter, dis_bytes, ct = F32(1.0), F32(-1.0), F32(1.0)
    status = 0
    for i in range(len(stack) - 1, -1, -1):
        cur = stack[i]
        find = any(var_in_expr(e, cur) for ix in indices for e in ix)
        if status == 0:
            if find:
                dis_iter = F32(dis_iter * F32(cur.extent))
                dis_bytes = F32(0.0)
                for accs in cur.touch.values():
                    for _, size, nbytes in accs:
                        dis_bytes = F32(dis_bytes + F32(size * nbytes))
            else:
                if dis_bytes < 0:
                    dis_bytes = F32(0.0)
                    for accs in cur.touch.values():
                        for _, _, nbytes in accs:
                            dis_bytes = F32(dis_bytes + F32(nbytes))
                status = 1
                ct = F32(ct * F32(cur.extent))
        else:
            if find:
                return R_LOOP_MULTIPLE_READ, dis_iter, dis_bytes, ct
            ct = F32(ct * F32(cur.extent))
        if len(cur.touch[buf]) - 1 > 0:                                   # SerialMultipleReadWrite, :568-596
            dis_iter = min(F32(size) for _, size, _ in cur.touch[buf])
            dis_bytes = F32(0.0)
            for accs in cur.touch.values():
                for acc_type, size, nbytes in accs:
                    if acc_type == K_READ:
                        dis_bytes = F32(dis_bytes + F32(size * nbytes))
            ct = F32(1.0)
            for j in range(i, -1, -1):
                ct = F32(ct * F32(stack[j].extent))
            return R_SERIAL_MULTIPLE_RW, F32(dis_iter / F32(cur.extent)), F32(dis_bytes / F32(cur.extent)), ct
    if status == 0:
        return R_NO_REUSE, F32(0), F32(0), F32(0)
    return R_LOOP_MULTIPLE_READ, dis_iter, dis_bytes, ct


def slog(x):
    """feature.cc:1051."""
    x = F32(x)
    return F32(-np.log2(F32(-x + F32(1)))) if x < 0 else F32(np.log2(F32(x + F32(1))))


class _Walker:
    """PerStoreFeatureExtractor, feature.cc:609-1048."""

    def __init__(self, template, S):
        self.t, self.S = template, S
        self.outer = F32(1.0)                     # outer_loop_prod_ (a float in the reference)
        self.stack, self.vec, self.unroll, self.par = [], [], [], []
        self.is_gpu = False
        self.tlen = {n: 1 for n in THREADS}
        self.auto_unroll = 0
        self.fea = {}                             # buffer id -> feature dict, in first-store order
        self.loops = {}                           # template loop node id -> _Loop (identity of the ForNode)

    def run(self, nodes):
        for n in nodes:
            kind = n["kind"]
            if kind in ("for", "vectorize", "unroll", "parallel"):
                self._for(n)
            elif kind == "thread":
                self._thread(n)
            elif kind == "pragma_unroll":
                old, self.auto_unroll = self.auto_unroll, self.S[n["value"]]
                self.run(n["body"])
                self.auto_unroll = old
            elif kind == "realize":
                self.run(n["body"])
                self._alloc(n)
            elif kind == "store":
                self._store(n)
            else:
                raise ValueError(kind)

    def _loop_of(self, n, kind):
        lp = self.loops.get(id(n))
        if lp is None:
            lp = self.loops[id(n)] = _Loop(n["name"], max(self.S[n["extent"]], 1), kind)
            n["_loop"] = lp
        return lp

    def _for(self, n):
        ext = max(self.S[n["extent"]], 1)               # an extent below 1 cannot come out of the snap; read as 1
        if ext == 1:                              # ScheduleOps drops trivial loops (MakeLoopNest, debug_keep_trivial_loop
            n["_loop"] = None                     # = false at feature.cc:1308): the var is its minimum, 0
            self.run(n["body"])
            return
        lp = self._loop_of(n, n["kind"])
        side = {"vectorize": self.vec, "unroll": self.unroll, "parallel": self.par}.get(n["kind"])
        if side is not None:
            side.append(lp)
        self.outer = F32(self.outer * F32(ext))
        self.stack.append(lp)
        self.run(n["body"])
        self.stack.pop()
        self.outer = F32(self.outer / F32(ext))
        if side is not None:
            side.pop()

    def _thread(self, n):
        """AttrStmt thread_extent / virtual_thread, feature.cc:613-658: a fake parallel For is pushed, even at extent 1."""
        name, ext = n["name"], max(self.S[n["extent"]], 1)
        before = self.tlen[name]
        self.tlen[name] = self.tlen[name] * ext if name == "vthread" else ext
        self.is_gpu = True
        lp = self._loop_of(n, "parallel")
        self.outer = F32(self.outer * F32(ext))
        self.stack.append(lp)
        self.run(n["body"])
        self.stack.pop()
        self.outer = F32(self.outer / F32(ext))
        self.tlen[name] = before

    def _resolve(self, e):
        """Template expressions name loops by node.  The var of a dropped (extent-1) loop is its minimum, 0, and the
        Simplify pass that precedes the extraction (feature.cc:1349-1351) folds it away: 0 * x = 0, 0 + x = x, x * 1 = x,
        0 // c = 0 % c = 0 -- the Mul / Add nodes the coefficient visitor reacts to disappear with it."""
        k = e[0]
        if k == "v":
            lp = e[1].get("_loop")
            return ("v", lp) if lp is not None else ("c", "zero")
        if k == "c":
            return e if self.S[e[1]] != 0 else ("c", "zero")
        a, b = self._resolve(e[1]), self._resolve(e[2])
        za, zb = a == ("c", "zero"), b == ("c", "zero")
        one = lambda x: x[0] == "c" and x[1] != "zero" and self.S[x[1]] == 1
        if k == "+":
            return b if za else a if zb else ("+", a, b)
        if k == "*":
            if za or zb:
                return ("c", "zero")
            return b if one(a) else a if one(b) else ("*", a, b)
        if za:
            return ("c", "zero")
        if k == "//" and one(b):
            return a
        if k == "%" and one(b):
            return ("c", "zero")
        # the rewrite simplifier knows every loop variable's range: a numerator that stays below the divisor folds
        Sd = dict(enumerate(self.S))
        Sd["zero"] = 0
        full = {id(lp): (0, lp.extent - 1) for lp in self.stack}
        hi, c = int_bound(a, full, Sd)[1], int_bound(b, full, Sd)[0]
        if hi < c:
            return ("c", "zero") if k == "//" else a
        return (k, a, b)

    def _store(self, n):
        S = self.S
        fea = self.fea.setdefault(n["buffer"], {"alloc_size": F32(0), "alloc_prod": F32(0), "alloc_outer_prod": F32(0),
                                                "alloc_inner_prod": F32(0)})
        ops = {k: int(n.get("ops", {}).get(k, 0)) for k in OPS}
        for _, idx in n.get("reads", []):
            for e in idx:
                index_ops(self._resolve(e), ops)
        # ---- group 1, :727-802
        for k in OPS:
            fea[k] = F32(self.outer * F32(ops[k]))
        for tag, st in (("vec", self.vec), ("unroll", self.unroll), ("parallel", self.par)):
            fea[tag + "_num"] = F32(len(st))
            fea[tag + "_len"], fea[tag + "_type"] = F32(0), POS_NONE
            if st:                                 # *_prod keeps its previous value otherwise (:743-777); zero-initialised
                fea[tag + "_len"] = F32(st[-1].extent)
                p = F32(1.0)
                for lp in st:
                    p = F32(p * F32(lp.extent))
                fea[tag + "_prod"], fea[tag + "_type"] = p, POS_MIXED
            fea.setdefault(tag + "_prod", F32(0))
        fea["is_gpu"] = F32(1.0 if self.is_gpu else 0.0)
        for name in THREADS:
            fea[name] = F32(self.tlen[name])
        # ---- accesses of this statement (BufferAccessExtractor, :342-379)
        acc = {}                                   # buffer id -> [acc_type, [index tuples]]
        acc[n["buffer"]] = [K_WRITE, [[self._resolve(e) for e in n["indices"]]]]
        for buf, idx in n.get("reads", []):
            a = acc.setdefault(buf, [K_UNKNOWN, []])
            if a[0] == K_WRITE:
                a[0] = K_READWRITE
            elif a[0] == K_UNKNOWN:
                a[0] = K_READ
            if a[0] != K_READWRITE:
                a[1].append([self._resolve(e) for e in idx])
        S = dict(enumerate(S))
        S["zero"] = 0
        # ---- touched regions of every enclosing loop, innermost first (:806-850)
        rng = {id(lp): (0, 0) for lp in self.stack}
        cur_ops = float(sum(ops[k] for k in OPS[:7]))
        mem_list, ops_list = [], []
        tmp_region = []
        bufs = self.t["buffers"]
        for i in range(len(self.stack) - 1, -1, -1):
            lp = self.stack[i]
            rng[id(lp)] = (0, lp.extent - 1)
            mem_bytes = 0
            for b, (acc_type, indices) in acc.items():
                tmp_region = compute_region(indices, rng, S)
                touched = 1
                for r in tmp_region:
                    touched *= r
                lp.touch.setdefault(b, []).append((acc_type, touched, bufs[b]["bytes"]))
                mem_bytes += touched * bufs[b]["bytes"]
            mem_list.append(F32(math.log2(mem_bytes)))
            cur_ops *= lp.extent
            ops_list.append(F32(math.log2(cur_ops)) if cur_ops > 0 else F32(-np.inf))
        # ---- per-buffer features (:853-940)
        feas = []
        for b, (acc_type, indices) in acc.items():
            shape, nb = bufs[b]["shape"], bufs[b]["bytes"]
            nbytes = F32(self.outer * F32(nb))
            if not self.stack:
                unique_bytes, stride, lines, unique_lines = F32(nb), 0, F32(1), F32(1)
            else:
                unique_bytes = F32(self.stack[0].touch[b][0][1] * nb)        # integer product, then float (:863-864)
                stride, reduce_ratio = 0, 1
                i = len(self.stack) - 1
                while i >= 0:
                    stride = compute_stride(indices, shape, self.stack[i], S)
                    if stride != 0:
                        break
                    reduce_ratio *= self.stack[-1].extent              # sic: always the innermost loop, :872
                    i -= 1
                lines = F32(F32(self.outer / F32(reduce_ratio)) * min(F32(1.0), F32(F32(1.0) * F32(stride * nb) / F32(CACHE_LINE))))
                lines = max(lines, F32(1.0))
                stride = stride if i == len(self.stack) - 1 else 0
                n_cont = F32(nb)
                for d in range(min(len(tmp_region), len(shape)) - 1, -1, -1):
                    if tmp_region[d] == shape[d]:
                        n_cont = F32(n_cont * F32(tmp_region[d]))
                        break
                unique_lines = max(F32(unique_bytes / min(n_cont, F32(CACHE_LINE))), F32(1.0))
            rtype, dis_iter, dis_bytes, ct = compute_reuse(b, indices, self.stack)
            f = {"acc_type": acc_type, "bytes": nbytes, "unique_bytes": unique_bytes, "lines": lines, "unique_lines": unique_lines,
                 "reuse_type": rtype, "reuse_dis_iter": dis_iter, "reuse_dis_bytes": dis_bytes, "reuse_ct": ct,
                 "stride": F32(stride), "buffer": b}
            if ct > 0.5:
                f.update(bytes_d=F32(nbytes / ct), unique_bytes_d=F32(unique_bytes / ct), lines_d=F32(lines / ct),
                         unique_lines_d=F32(unique_lines / ct))
            else:
                f.update(bytes_d=F32(nbytes * 2), unique_bytes_d=F32(unique_bytes * 2), lines_d=F32(lines * 2),
                         unique_lines_d=F32(unique_lines * 2))
            feas.append(f)
        fea["access"] = feas
        # ---- group 3: arithmetic intensity curve (:943-979)
        curve = [F32(0)] * ARITH_N
        if cur_ops > 0 and ops_list:
            pt = 0
            for i in range(ARITH_N):
                x = F32(ops_list[-1] * F32(i + 1) / F32(ARITH_N))
                while float(ops_list[pt]) < float(x) - 1e-4:                 # float - double literal: compared in double (:958)
                    pt += 1
                if pt == 0:
                    v = F32(ops_list[0] / mem_list[0])
                else:
                    base = F32(ops_list[pt - 1] / mem_list[pt - 1])
                    slope = F32(F32(F32(ops_list[pt] / mem_list[pt]) - base) / F32(ops_list[pt] - ops_list[pt - 1]))
                    v = F32(base + slope * F32(x - ops_list[pt - 1]))
                curve[i] = v
        fea["curve"] = curve
        # ---- group 5 (:997-1003)
        fea["outer_prod"], fea["num_loops"], fea["auto_unroll_max_step"] = self.outer, F32(len(self.stack)), F32(self.auto_unroll)

    def _alloc(self, n):
        """ExtractAllocationFeature, feature.cc:982-994 (called after the body of the BufferRealize)."""
        fea = self.fea.setdefault(n["buffer"], {})
        size = F32(1.0)
        for s in n["bounds"]:
            size = F32(size * F32(self.S[s]))
        fea["alloc_size"] = F32(size * F32(self.t["buffers"][n["buffer"]]["bytes"]))
        fea["alloc_prod"] = F32(size * self.outer)
        fea["alloc_outer_prod"] = self.outer
        fea["alloc_inner_prod"] = F32(fea.get("outer_prod", F32(0)) / self.outer)


def _strip(nodes):
    for n in nodes:
        n.pop("_loop", None)
        if "body" in n:
            _strip(n["body"])


def feature_sets(template, extents):
    """The FeatureSet of every stored buffer, as dicts, in first-store order."""
    S = scalar_values(template, extents)
    w = _Walker(template, S)
    try:
        w.run(template["body"])
    finally:
        _strip(template["body"])
    return {b: f for b, f in w.fea.items() if "outer_prod" in f}


def vectorise(fea):
    """One 164-d row from a FeatureSet: GetPerStoreFeature, feature.cc:1060-1187."""
    r = [slog(fea[k]) for k in OPS]
    for tag in ("vec", "unroll", "parallel"):
        r += [slog(fea[tag + "_num"]), slog(fea[tag + "_prod"]), slog(fea[tag + "_len"])]
        r += [F32(1.0 if i == fea[tag + "_type"] else 0.0) for i in range(8)]
    r.append(fea["is_gpu"])
    r += [slog(fea[n]) for n in THREADS]
    acc = fea["access"]
    order = sorted(range(len(acc)), key=lambda i: (float(acc[i]["reuse_dis_bytes"]), float(acc[i]["unique_bytes"]),
                                                   float(acc[i]["unique_lines"]), float(acc[i]["acc_type"])))[:MAX_N_BUFS]
    for i in order:
        a = acc[i]
        r += [F32(1.0 if j == a["acc_type"] else 0.0) for j in range(3)]
        r += [slog(a["bytes"]), slog(a["unique_bytes"]), slog(a["lines"]), slog(a["unique_lines"])]
        r += [F32(1.0 if j == a["reuse_type"] else 0.0) for j in range(3)]
        r += [slog(a["reuse_dis_iter"]), slog(a["reuse_dis_bytes"]), slog(a["reuse_ct"]), slog(a["bytes_d"]),
              slog(a["unique_bytes_d"]), slog(a["lines_d"]), slog(a["unique_lines_d"]), slog(a["stride"])]
    r += [F32(0)] * (18 * (MAX_N_BUFS - len(order)))
    r += list(fea["curve"])
    r += [slog(fea["alloc_size"]), slog(fea["alloc_prod"]), slog(fea["alloc_outer_prod"]), slog(fea["alloc_inner_prod"])]
    r += [slog(fea["outer_prod"]), slog(fea["num_loops"]), slog(fea["auto_unroll_max_step"])]
    assert len(r) == ROW
    return np.array(r, dtype=np.float32)


def evaluate(template, extents):
    """float32 [stored buffers, 164] (first-store order) and the buffer id of every row."""
    fs = feature_sets(template, extents)
    rows = np.stack([vectorise(f) for f in fs.values()]) if fs else np.zeros((0, ROW), np.float32)
    return rows, list(fs.keys())


def feature_names():
    """GetPerStoreFeatureName, feature.cc:1190-1293 (the names the reference's tests index by)."""
    n = ["float_mad", "float_addsub", "float_mul", "float_divmod", "float_cmp", "float_mathfunc", "float_otherfunc",
         "int_mad", "int_addsub", "int_mul", "int_divmod", "int_cmp", "int_mathfunc", "int_otherfunc", "bool_op", "select_op"]
    pos = ["kPosNone", "kPosInnerSpatial", "kPosMiddleSpatial", "kPosOuterSpatial", "kPosInnerReduce", "kPosMiddleReduce",
           "kPosOuterReduce", "kPosMixed"]
    for tag in ("vec", "unroll", "parallel"):
        n += [tag + "_num", tag + "_prod", tag + "_len"] + [tag + "_type." + p for p in pos]
    n += ["is_gpu", "blockIdx_x_len", "blockIdx_y_len", "blockIdx_z_len", "threadIdx_x_len", "threadIdx_y_len",
          "threadIdx_z_len", "vthread_len"]
    for i in range(MAX_N_BUFS):
        p = "B%d." % i
        n += [p + "acc_type.kRead", p + "acc_type.kWrite", p + "acc_type.kReadWrite", p + "bytes", p + "unique_bytes",
              p + "lines", p + "unique_lines", p + "reuse_type.kLoopMultipleRead", p + "reuse_type.kSerialMultipleReadWrite",
              p + "reuse_type.kNoReuse", p + "reuse_dis_iter", p + "reuse_dis_bytes", p + "reuse_ct", p + "bytes_d_reuse_ct",
              p + "unique_bytes_d_reuse_ct", p + "lines_d_reuse_ct", p + "unique_lines_d_reuse_ct", p + "stride"]
    n += ["arith_intensity_curve_%d" % i for i in range(ARITH_N)]
    n += ["alloc_size", "alloc_prod", "alloc_outer_prod", "alloc_inner_prod", "outer_prod", "num_loops", "auto_unroll_max_step"]
    assert len(n) == ROW
    return n
