"""Extent parser, scaler and the way back to measure records.

Mirror of scripts/pre_experiments/model_myself/utils/extent.py (find_common_for_loops :8-30,
remove_common_for_loops :33-88, state_to_records :94-136, gen_program_pool :140-174) and of the scaler
cell of the notebook (vae_extent_search.py:394-397: np.log1p + sklearn StandardScaler).

* the numeric part of state_to_records -- two regular expressions per printed state -- is one pass over the
  bytes of all states in libvaex_b200.so (host C++, all host threads), straight into a pinned fp32 matrix;
* ExtentScaler.fit / transform run on the GPU (fp64 column reductions with sklearn's variance and
  constant-column rules, sklearn's fp32 arithmetic in transform); inverse_transform is the snap kernel;
* records_from_extents closes the loop the reference left open: decoded, legalised extents are written back
  into the SplitStep lengths / auto_unroll pragma of a sketch's measure record
  (src/auto_scheduler/transform_step.cc:1015-1023,657-664; record keys "i"/"r"/"v", measure_record.cc:320-340).
"""
import ctypes
import json
import re

import numpy as np

from . import _lib

_TRIVIAL_FOR = re.compile(r"for\s+(\S+)\s+\(0,\s*1\)")


# ------------------------------------------------------------------ text clean-up (display only, host)
def find_common_for_loops(schedules):
    """Names of the `for v (0,1)` loops that occur in every schedule (utils/extent.py:8-30)."""
    common = None
    for sched in schedules:
        names = {m.group(1) for m in (_TRIVIAL_FOR.match(ln.lstrip()) for ln in sched.split("\n")) if m}
        common = names if common is None else common & names
    return common if common is not None else set()


def remove_common_for_loops(schedule, common_vars):
    """Drops the trivial loops named in common_vars and pulls their bodies two columns left
    (utils/extent.py:33-88).  One pass with a stack of the open dropped loops: a dropped loop stays open until
    the first kept, non-blank line indented no deeper than the loop itself."""
    out, open_loops = [], []
    for line in schedule.split("\n"):
        body = line.lstrip()
        indent = len(line) - len(body)
        m = _TRIVIAL_FOR.match(body)
        if m and m.group(1) in common_vars:
            open_loops.append(indent)
            continue
        if not line.strip():
            out.append(line)
            continue
        open_loops = [b for b in open_loops if b < indent]
        out.append(" " * max(0, indent - 2 * len(open_loops)) + body)
    return "\n".join(out)


# ------------------------------------------------------------------ parser
def _pinned_empty(shape, dtype):
    """Pinned numpy array when torch + CUDA are there (async H2D), plain numpy otherwise."""
    try:
        import torch
        if torch.cuda.is_available():
            t = torch.empty(shape, dtype=getattr(torch, np.dtype(dtype).name), pin_memory=True)
            return t.numpy(), t
    except Exception:
        pass
    return np.empty(shape, dtype=dtype), None


def parse_extents(state_strs, d=None):
    """Printed states -> (matrix fp32 [n, d], row_len int32 [n], counts int32 [n, 2]).  d defaults to the longest
    row.  Rows shorter than d are zero-padded; a row longer than an explicit d raises."""
    lib = _lib.load()
    enc = [s.encode("utf-8") if isinstance(s, str) else bytes(s) for s in state_strs]
    n = len(enc)
    offs = np.zeros(n + 1, dtype=np.int64)
    if n:
        np.cumsum([len(b) for b in enc], out=offs[1:])
    text = b"".join(enc)
    buf = ctypes.create_string_buffer(text, len(text) + 1)
    counts = np.zeros((n, 2), dtype=np.int32)
    _lib.check(lib.vaex_extent_text_scan(ctypes.addressof(buf), offs.ctypes.data, n, counts.ctypes.data))
    need = int(counts.sum(axis=1).max()) if n else 0
    d = max(need, 1) if d is None else int(d)
    mat, keep = _pinned_empty((n, d), np.float32)
    row_len = np.zeros(n, dtype=np.int32)
    _lib.check(lib.vaex_extent_text_parse(ctypes.addressof(buf), offs.ctypes.data, n, d, mat.ctypes.data, row_len.ctypes.data))
    if keep is not None:
        mat = _Pinned(mat, keep)
    return mat, row_len, counts


class _Pinned(np.ndarray):
    """ndarray view that keeps the pinned torch tensor it aliases alive."""

    def __new__(cls, arr, owner):
        obj = arr.view(cls)
        obj._owner = owner
        return obj

    def __array_finalize__(self, obj):
        self._owner = getattr(obj, "_owner", None)


def state_to_records(state_strs, verbose=False):
    """utils/extent.py:94-136.  Same keys ("schedules", "extents", "unroll", "all", "cleaned_schedules") plus
    "matrix": the [n, d] fp32 matrix (pinned when CUDA is present) that np.array(records["all"]) would build."""
    states = [s.strip() for s in state_strs]
    records = {"schedules": [st.split("Placeholder")[-1][2:] for st in states]}
    mat, row_len, counts = parse_extents(states)
    records["extents"] = [[float(v) for v in mat[i, :counts[i, 0]]] for i in range(len(states))]
    records["unroll"] = [[float(v) for v in mat[i, counts[i, 0]:row_len[i]]] for i in range(len(states))]
    records["all"] = [np.array(mat[i, :row_len[i]], dtype=np.float32) for i in range(len(states))]
    records["matrix"] = mat
    common = find_common_for_loops(records["schedules"])
    records["cleaned_schedules"] = [remove_common_for_loops(s, common) for s in records["schedules"]]
    if verbose:
        print("common (0,1) loops:", common)
    return records


def gen_program_pool(task, size, evo_population, min_population, seed=2023):
    """utils/extent.py:140-174 drives TVM's SketchPolicy (sample_initial_population + evolutionary_search);
    that is Ansor's C++ search, outside this library.  With a real tvm importable the same calls are made and
    the printed states go through state_to_records; without it this raises."""
    try:
        from tvm import auto_scheduler
        policy = auto_scheduler.SketchPolicy(
            task, auto_scheduler.XGBModel(),
            params={"evolutionary_search_num_iters": 4, "evolutionary_search_population": evo_population,
                    "sample_init_min_population": min_population}, seed=seed, verbose=0)
    except Exception as e:  # no tvm (or the test stub)
        raise _lib.VaexError("gen_program_pool needs TVM's SketchPolicy (not part of vaex_b200): %r" % (e,))
    states = policy.evolutionary_search(policy.sample_initial_population(), size * 2)
    seen = []
    for s in states:
        t = str(s)
        if t not in seen:
            seen.append(t)
            if len(seen) >= size:
                break
    return state_to_records(seen)


# ------------------------------------------------------------------ scaler
class ExtentScaler:
    """np.log1p followed by sklearn's StandardScaler (vae_extent_search.py:394-397), on the GPU.
    fit / transform take RAW extents (fp32 [n, d], host or CUDA); mean_, var_, scale_ are float64 numpy arrays
    like sklearn's.  inverse_transform gives the integer extents back: clamp(round(expm1(x*scale_+mean_)), 1)."""

    _SCRATCH_BLOCKS = 592

    def __init__(self):
        self.mean_ = self.var_ = self.scale_ = None
        self.n_samples_seen_ = 0
        self._dev = None

    @staticmethod
    def _to_cuda(x):
        import torch
        if isinstance(x, np.ndarray):
            t = torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32))
            return t.cuda(non_blocking=True)
        if not x.is_cuda:
            x = x.cuda()
        return x.detach().contiguous().float()

    def fit(self, raw):
        import torch
        x = self._to_cuda(raw)
        n, d = x.shape
        stats = torch.empty(3, d, dtype=torch.float64, device=x.device)
        scratch = torch.empty(2 * d * self._SCRATCH_BLOCKS, dtype=torch.float64, device=x.device)
        _lib.check(_lib.load().vaex_scaler_fit(_lib.ptr(x), n, d, _lib.ptr(stats[0]), _lib.ptr(stats[1]), _lib.ptr(stats[2]),
                                              _lib.ptr(scratch), scratch.numel(), _lib.stream_ptr()))
        self._dev = stats
        host = stats.cpu().numpy()
        self.mean_, self.var_, self.scale_ = host[0].copy(), host[1].copy(), host[2].copy()
        self.n_samples_seen_ = n
        return self

    def _stats(self, device):
        import torch
        if self.mean_ is None:
            raise _lib.VaexError("ExtentScaler is not fitted")
        if self._dev is None or self._dev.device != device:
            self._dev = torch.from_numpy(np.stack([self.mean_, np.zeros_like(self.mean_) if self.var_ is None else self.var_,
                                                   self.scale_])).to(device)
        return self._dev

    def transform(self, raw):
        """-> CUDA fp32 [n, d] = StandardScaler.transform(np.log1p(raw))."""
        import torch
        x = self._to_cuda(raw)
        st = self._stats(x.device)
        out = torch.empty_like(x)
        _lib.check(_lib.load().vaex_scaler_transform(_lib.ptr(x), x.shape[0], x.shape[1], _lib.ptr(st[0]), _lib.ptr(st[2]),
                                                    _lib.ptr(out), _lib.stream_ptr()))
        return out

    def fit_transform(self, raw):
        x = self._to_cuda(raw)
        return self.fit(x).transform(x)

    def inverse_transform(self, x_scaled, slot_extent=None, slot_last=None, max_innermost=64):
        """scaled fp32 [n, d] -> int32 extents (snap kernel, VAEX_SNAP_EXPM1; divisor-legal with slot tables)."""
        import torch
        from .vae_model import snap_extents
        x = self._to_cuda(x_scaled)
        mean = torch.from_numpy(self.mean_.astype(np.float32)).to(x.device)
        scale = torch.from_numpy(self.scale_.astype(np.float32)).to(x.device)
        return snap_extents(x, mean, scale, log1p=True, slot_extent=slot_extent, slot_last=slot_last, max_innermost=max_innermost)


# ------------------------------------------------------------------ extents -> measure records
class SketchSlots:
    """Which transform step each slot of the extent vector writes, for ONE sketch.

    slots[j] is ("SP", step_index, length_index) -- the j-th printed extent is lengths[length_index] of that
    SplitStep; ("PR", step_index) -- the auto_unroll_max_step pragma of that PragmaStep; or None -- a derived
    value (the outermost part extent / prod(lengths), a loop that is not split, an inlined constant) that is
    recomputed rather than written.  The reference never built this map; it comes from the
    sketch (Ansor's sketch rules fix which iterator every tile level of every stage prints as)."""

    def __init__(self, template_record, slots):
        self.template = json.loads(template_record) if isinstance(template_record, (str, bytes)) else template_record
        self.steps = self.template["i"][1][1]
        self.slots = list(slots)
        for j, s in enumerate(self.slots):
            if s is None:
                continue
            step = self.steps[s[1]]
            if s[0] == "SP":
                if step[0] != "SP" or not (0 <= s[2] < len(step[4])):
                    raise ValueError("slot %d does not name a SplitStep length: %r" % (j, step))
            elif s[0] == "PR":
                if step[0] != "PR" or not str(step[3]).startswith("auto_unroll_max_step"):
                    raise ValueError("slot %d does not name an auto_unroll pragma: %r" % (j, step))
            else:
                raise ValueError("unknown slot kind %r" % (s[0],))

    @property
    def d(self):
        return len(self.slots)

    def legalise_tables(self):
        """(slot_extent int32[d], slot_last uint8[d], slot_order int32[m]) for vaex_legalise_extents: the lengths of
        one SplitStep consecutive in walk order (innermost level last), the step's extent on each of them."""
        slot_extent = np.zeros(self.d, dtype=np.int32)
        slot_last = np.zeros(self.d, dtype=np.uint8)
        groups = {}
        for j, s in enumerate(self.slots):
            if s is not None and s[0] == "SP":
                groups.setdefault(s[1], []).append((s[2], j))
        order = []
        for step_index in sorted(groups):
            members = sorted(groups[step_index])
            ext = int(self.steps[step_index][3])
            for _, j in members:
                slot_extent[j] = ext
                order.append(j)
            slot_last[members[-1][1]] = 1
        return slot_extent, slot_last, np.asarray(order, dtype=np.int32)

    def legalise(self, extents, max_innermost=0):
        """In place on a CUDA int32 [n, d] tensor: every SplitStep's lengths multiply to a divisor of its extent."""
        import torch
        if not (extents.is_cuda and extents.dtype == torch.int32 and extents.is_contiguous()):
            raise _lib.VaexError("legalise needs a contiguous CUDA int32 tensor")
        se, sl, so = (torch.from_numpy(a).to(extents.device) for a in self.legalise_tables())
        _lib.check(_lib.load().vaex_legalise_extents(_lib.ptr(extents), extents.shape[0], extents.shape[1], _lib.ptr(se), _lib.ptr(sl),
                                                    _lib.ptr(so), so.numel(), int(max_innermost), _lib.stream_ptr()))
        return extents

    def records_from_extents(self, extents, strict=True):
        """int extents [n, d] (numpy) -> list of record JSON lines (same "i"/"r"/"v" layout as the template, "r" reset
        to an unmeasured result).  strict: raise when a SplitStep's lengths do not multiply to a divisor of its extent
        (Ansor would reject the state); otherwise such rows come back as None."""
        ext = np.asarray(extents)
        out = []
        for row in ext:
            steps = json.loads(json.dumps(self.steps))
            for j, s in enumerate(self.slots):
                if s is None:
                    continue
                if s[0] == "SP":
                    steps[s[1]][4][s[2]] = int(row[j])
                else:
                    steps[s[1]][3] = "auto_unroll_max_step$%d" % int(row[j])
            bad = None
            for k, st in enumerate(steps):
                if st[0] == "SP" and int(st[3]) > 0:
                    prod = 1
                    for ln in st[4]:
                        prod *= int(ln)
                    if prod <= 0 or int(st[3]) % prod:
                        bad = (k, st)
                        break
            if bad is not None:
                if strict:
                    raise ValueError("step %d: lengths %r do not divide extent %d" % (bad[0], bad[1][4], bad[1][3]))
                out.append(None)
                continue
            rec = {"i": [self.template["i"][0], [self.template["i"][1][0], steps]],
                   "r": [[0], 0, 0, 0], "v": self.template.get("v", "v0.6")}
            out.append(json.dumps(rec, separators=(", ", ": ")))
        return out

    def extents_from_record(self, record):
        """The written slots of a record (derived slots = 0): the inverse of records_from_extents."""
        rec = json.loads(record) if isinstance(record, (str, bytes)) else record
        steps = rec["i"][1][1]
        row = np.zeros(self.d, dtype=np.int32)
        for j, s in enumerate(self.slots):
            if s is None:
                continue
            row[j] = int(steps[s[1]][4][s[2]]) if s[0] == "SP" else int(str(steps[s[1]][3]).split("$")[1])
        return row
