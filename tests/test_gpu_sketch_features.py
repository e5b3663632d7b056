"""n4: the CUDA evaluator of sketch templates (csrc/sketch_features.cu) against the restated PerStoreFeatureExtractor
(oracle/perstore_features.py, pinned to the reference's known-answer tests in tests/test_oracle_perstore.py) on the same
templates and the same random extents.  Counts, one-hot fields and integer-valued features must agree exactly; log2-valued
ones within 2e-6 relative (device log2f against numpy's)."""
import numpy as np
import pytest
import torch

from oracle import perstore_features as PF
from vaex_b200 import _lib, sketch_features as SF
from test_oracle_perstore import fusion_template, matmul_gpu_template

pytestmark = pytest.mark.gpu


def check(template, extents):
    sk = SF.SketchFeatures(template)
    got = sk.rows(torch.from_numpy(np.asarray(extents, np.int32)).cuda()).cpu().numpy().reshape(len(extents), sk.n_rows, 164)
    td = template.as_dict()
    worst = 0.0
    for i, e in enumerate(extents):
        want, bufs = PF.evaluate(td, list(e))
        assert bufs == template.stored_buffers()
        assert np.isfinite(got[i]).all()
        err = np.abs(got[i] - want) / np.maximum(1.0, np.abs(want))
        worst = max(worst, float(err.max()))
        assert err.max() < 2e-6, (i, list(e), np.argwhere(err > 2e-6)[:8].tolist(), got[i][err > 2e-6][:8], want[err > 2e-6][:8])
    return worst


def test_known_answer_templates():
    check(SF.matmul_cpu_template(512, 512, 512), [[32, 64, 512, 8, 16]])
    check(fusion_template(64, 32), [[64, 32]])
    check(matmul_gpu_template(), [[1]])


def test_matmul_template_random_extents():
    rng = np.random.default_rng(3)
    div = np.array([1, 1, 2, 3, 4, 7, 8, 16, 32, 64, 128, 512])
    ext = rng.choice(div, size=(300, 5))
    ext[0] = 1                                   # every loop trivial: no loop nest at all
    ext[1] = [1, 1, 512, 1, 1]
    check(SF.matmul_cpu_template(512, 512, 512), ext.tolist())


def test_conv2d_template_random_extents():
    t = SF.conv2d_nhwc_cpu_template()
    rng = np.random.default_rng(4)
    div = np.array([1, 1, 2, 3, 4, 7, 8, 14, 16, 28])
    ext = rng.choice(div, size=(200, t.n_slots))
    ext[0] = SF.CONV2D_README_EXTENTS
    check(t, ext.tolist())


def test_record_derived_conv2d_template():
    """The template and extents read off the reference's measured ci_logs record (padding stage attached under ax3.0 and
    vectorised, 7 x 7 stride-2 conv on 224 x 224 x 3), plus random tilings of it."""
    from test_cpu_sketch_template import CI_LOG_STEPS
    key = '["12b88bedece6984af589a28b43e0f3c4", 1, 224, 224, 3, 7, 7, 3, 64, 1, 1, 1, 64, 1, 112, 112, 64]'
    t, ext = SF.conv2d_template_from_record({"i": [[key, "llvm"], [[], CI_LOG_STEPS]]})
    rng = np.random.default_rng(8)
    more = rng.choice(np.array([1, 1, 2, 4, 7, 8, 16]), size=(120, t.n_slots))
    check(t, [ext] + more.tolist())


def test_matmul_multi_level_tiling_template_random_extents():
    rng = np.random.default_rng(6)
    div = np.array([1, 1, 2, 4, 8, 16, 32, 64])
    for derive in (False, True):
        t = SF.matmul_mlt_cpu_template(512, 512, 512, derive_outer=derive)
        ext = rng.choice(div, size=(150, t.n_slots))
        if derive:                                # the derived outer level needs a legal tiling: inner levels multiply to a divisor of 512
            ext = rng.choice(np.array([1, 2, 4, 8]), size=(150, t.n_slots))
        ext[:, 10] = rng.choice([0, 16, 64, 512], size=150)
        check(t, ext.tolist())


def test_ragged_batch_and_errors():
    t = SF.matmul_cpu_template(64, 64, 64)
    sk = SF.SketchFeatures(t)
    for n in (1, 31, 33, 64, 65, 1000):
        e = torch.randint(1, 9, (n, 5), dtype=torch.int32, device="cuda")
        rows = sk.rows(e)
        assert rows.shape == (n * sk.n_rows, 164) and torch.isfinite(rows).all()
        again = sk.rows(e[n // 2:n // 2 + 1])
        assert torch.equal(again, rows[(n // 2) * sk.n_rows:(n // 2 + 1) * sk.n_rows])     # a row depends on its candidate only
    assert sk.rows(torch.empty(0, 5, dtype=torch.int32, device="cuda")).shape == (0, 164)
    with pytest.raises(ValueError):
        sk.rows(torch.ones(4, 6, dtype=torch.int32, device="cuda"))
    with pytest.raises(_lib.VaexError):
        sk.rows(torch.ones(4, 5, dtype=torch.int32))
    bad = sk.program.clone()
    bad[0] = 7
    out = torch.empty(4, 164, device="cuda")
    with pytest.raises(_lib.VaexError, match="not a sketch program"):
        _lib.check(_lib.load().vaex_sketch_features(_lib.ptr(bad), bad.numel(), _lib.ptr(torch.ones(4, 5, dtype=torch.int32, device="cuda")),
                                                   4, _lib.ptr(out), _lib.stream_ptr()))


def test_sketch_scorer_pipeline():
    """z -> decode + snap + legalise -> feature rows -> SegmentSum-MLP, device-resident: the extents are legal tile factors of
    the conv2d task, the rows are the oracle's for those extents, and the scores are the cost model's on those rows (the
    reference order of operations: decode, lower + extract on the CPU, MLPModel.predict)."""
    import vaex_b200
    from oracle import torch_oracle as O
    torch.manual_seed(5)
    t = SF.conv2d_nhwc_cpu_template(derive_outer=True)
    se, sl = SF.conv2d_slot_tables()
    vae = vaex_b200.VAE_feature_head(23, None, 64, 256).cuda().eval()
    net = vaex_b200.SegmentSumMLPModule(164, 256, 1).cuda().eval()
    rng = np.random.default_rng(5)
    mean, scale = rng.uniform(0.5, 3, 23).astype(np.float32), rng.uniform(0.3, 1.5, 23).astype(np.float32)
    sc = SF.SketchScorer(vae, mean, scale, t, net, se, sl)
    n = 777
    z = torch.randn(n, 64, device="cuda")
    scores, ext = sc.score(z)
    e = ext.cpu().numpy()
    assert e.shape == (n, 23) and (e >= 1).all()
    for key, E in (("y", 28), ("x", 28), ("f", 128), ("ry", 3), ("rx", 3), ("rc", 64)):
        inner = e[:, list(SF.CONV2D_SLOTS[key][1:])].astype(np.int64).prod(1)
        assert (E % inner == 0).all(), key                                   # SplitStep lengths multiply to a divisor of the axis
    td = t.as_dict()
    pick = rng.choice(n, 40, replace=False)
    rows = sc.features.rows(ext).cpu().numpy().reshape(n, 3, 164)
    for i in pick:
        want, _ = PF.evaluate(td, e[i].tolist())
        assert np.abs(rows[i] - want).max() < 2e-6 * max(1.0, np.abs(want).max())
    P = {k: v.detach().cpu() for k, v in net.state_dict().items()}
    want = O.segsum_mlp_forward(P, np.full(n, 3, np.int32), torch.from_numpy(rows.reshape(-1, 164))).numpy()
    got = scores.cpu().numpy()
    assert np.linalg.norm(got - want) / np.linalg.norm(want) < 1e-3
    again, _ = sc.score(z)
    assert torch.equal(again, scores)


def test_rows_at_bench_size_are_position_independent():
    """Size-independent properties at the bench's batch (2^20 candidates would need 2 GB twice; 2^18 here): every value is
    finite, a candidate's rows do not depend on where it sits in the batch (a permutation of the extents permutes the rows
    bit for bit), and the work-invariant features agree across all candidates of the task (float ops of the Conv2dOutput row:
    57.8 M multiplies whatever the tiling)."""
    t = SF.conv2d_nhwc_cpu_template(derive_outer=True)
    sk = SF.SketchFeatures(t)
    n = 1 << 18
    g = torch.Generator(device="cuda").manual_seed(9)
    e = torch.ones(n, 23, dtype=torch.int32, device="cuda")
    divs = {"y": [1, 2, 4, 7, 14, 28], "x": [1, 2, 4, 7, 14, 28], "f": [1, 2, 4, 8, 16, 32], "rc": [1, 2, 4, 8, 16, 64], "ry": [1, 3], "rx": [1, 3]}
    for key, ds in divs.items():
        sl = SF.CONV2D_SLOTS[key]
        d = torch.tensor(ds, dtype=torch.int32, device="cuda")
        e[:, sl[-1]] = d[torch.randint(0, len(ds), (n,), device="cuda", generator=g)]          # one non-trivial inner level per axis
    e[:, 22] = torch.tensor([0, 16, 64, 512], dtype=torch.int32, device="cuda")[torch.randint(0, 4, (n,), device="cuda", generator=g)]
    rows = sk.rows(e).view(n, 3, 164)
    assert torch.isfinite(rows).all()
    perm = torch.randperm(n, device="cuda", generator=g)
    assert torch.equal(sk.rows(e[perm].contiguous()).view(n, 3, 164), rows[perm])
    conv = rows[:, 1, :]
    want = float(np.log2(28 * 28 * 128 * 3 * 3 * 64 + 1.0))
    assert float((conv[:, 2] - want).abs().max()) < 1e-5 and float((conv[:, 1] - want).abs().max()) < 1e-5     # float_mul, float_addsub


def test_sketch_chain_is_graph_capturable():
    """After one warm-up call the whole chain (decode + snap + legalise -> feature rows -> SegmentSum-MLP) launches without
    touching the host: captured once into a CUDA graph, a replay on new latents reproduces the eager scores bit for bit."""
    import vaex_b200
    torch.manual_seed(11)
    t = SF.conv2d_nhwc_cpu_template(derive_outer=True)
    se, sl = SF.conv2d_slot_tables()
    vae = vaex_b200.VAE_feature_head(23, None, 64, 256).cuda().eval()
    net = vaex_b200.SegmentSumMLPModule(164, 256, 1).cuda().eval()
    rng = np.random.default_rng(11)
    sc = SF.SketchScorer(vae, rng.uniform(0.5, 3, 23).astype(np.float32), rng.uniform(0.3, 1.5, 23).astype(np.float32), t, net, se, sl)
    z = torch.randn(4096, 64, device="cuda")
    sc.score(z)                                            # warm-up: packs the models, checks the program, sizes the workspaces
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=side, capture_error_mode="thread_local"):
        out, _ = sc.score(z)
    torch.cuda.current_stream().wait_stream(side)
    z.copy_(torch.randn(4096, 64, device="cuda"))
    g.replay()
    torch.cuda.synchronize()
    replayed = out.clone()
    eager, _ = sc.score(z)
    assert torch.equal(replayed, eager)


def test_corrupted_programs_are_rejected():
    """The library checks every offset, count and id of a program the first time it sees it: a damaged program is an error
    return, not a kernel walking off its tables."""
    t = SF.conv2d_nhwc_cpu_template()
    good = t.compile()
    e = torch.ones(4, 23, dtype=torch.int32, device="cuda")
    out = torch.empty(12, 164, device="cuda")
    lib = _lib.load()
    o_in, o_st, o_ex = int(good[9]), int(good[10]), int(good[11])
    cases = {"token count": (o_ex, 1000), "instruction": (o_in, 99), "loop id": (o_in + 1, 5), "store buffer": (o_st, 77),
             "section order": (9, int(good[10]) + 5), "slots": (1, 0)}
    for name, (pos, val) in cases.items():
        bad = good.copy()
        bad[pos] = val
        prog = torch.from_numpy(bad).cuda()
        with pytest.raises(_lib.VaexError, match="sketch program"):
            _lib.check(lib.vaex_sketch_features(_lib.ptr(prog), prog.numel(), _lib.ptr(e), 4, _lib.ptr(out), _lib.stream_ptr()))
    prog = torch.from_numpy(good).cuda()
    _lib.check(lib.vaex_sketch_features(_lib.ptr(prog), prog.numel(), _lib.ptr(e), 4, _lib.ptr(out), _lib.stream_ptr()))
    assert torch.isfinite(out).all()


def test_mlpmodel_predict_extents_matches_predict_on_the_same_rows():
    """MLPModel.predict_extents(task, template, extents) = MLPModel.predict's scoring (mlp_model.py:832-846) on the rows the
    template implies: equal, at the tensor path's tolerance, to the reference order of operations -- rows for every
    candidate (here the oracle's), normalised by fea_norm_vec, SegmentSumMLP forward -- with and without the task embedding."""
    import vaex_b200
    from oracle import torch_oracle as O

    class _Task:
        workload_key, target = "dense-softmax-key", "llvm"

    t = SF.matmul_mlt_cpu_template(256, 256, 256)
    rng = np.random.default_rng(21)
    ext = rng.choice([1, 2, 4, 8, 16], size=(300, 11)).astype(np.int32)
    rows = np.concatenate([PF.evaluate(t.as_dict(), e.tolist())[0] for e in ext], 0)
    for emb_on in (False, True):
        torch.manual_seed(6)
        mm = vaex_b200.MLPModel()
        mm.model = vaex_b200.MLPModelInternal(device="cuda:0", use_workload_embedding=emb_on)
        F = 174 if emb_on else 164
        mm.model.base_model = vaex_b200.SegmentSumMLPModule(F, 256, 1).cuda()
        norm = np.concatenate([np.maximum(rows.max(0), 1e-3), np.full(F - 164, 1.0, np.float32)]).astype(np.float32)
        mm.model.fea_norm_vec = torch.from_numpy(norm)
        got = mm.predict_extents(_Task, t, ext)
        assert got.shape == (300,) and np.isfinite(got).all()
        emb = mm.model._task_embedding(vaex_b200.LearningTask(_Task.workload_key, _Task.target))
        full = rows if emb is None else np.concatenate([rows, np.broadcast_to(emb, (len(rows), 10))], 1)
        P = {k: v.detach().cpu() for k, v in mm.model.base_model.state_dict().items()}
        want = O.segsum_mlp_forward(P, np.full(300, t.n_rows, np.int32), torch.from_numpy((full / norm).astype(np.float32))).numpy()
        # fp16 operands on two-row programs with scores around 0.03: 1.6e-3 measured; the fp32 kernels pin the arithmetic
        assert np.linalg.norm(got - want) / np.linalg.norm(want) < 3e-3
        simt = mm.predict_extents(_Task, t, ext, flags=_lib.F_MASK_INVALID | _lib.F_FORCE_SIMT)
        assert np.linalg.norm(simt - want) / np.linalg.norm(want) < 2e-5
    mm = vaex_b200.MLPModel()
    mm.model = None
    assert mm.predict_extents(_Task, t, ext).shape == (300,)              # no trained model: random scores, like predict


def test_fp16_rows_are_the_narrowed_fp32_rows_and_score_identically():
    """vaex_sketch_features_f16 = cvt.rn.satfinite of the fp32 rows, element for element (odd batch sizes, the row tail of
    36 values, the allocation features patched in at the end); and because the tcgen05 row encoder narrows fp32 rows in the
    same way before its first GEMM, the SegmentSum-MLP scores from either are bit-identical."""
    import vaex_b200
    torch.manual_seed(13)
    net = vaex_b200.SegmentSumMLPModule(164, 256, 1).cuda().eval()
    for tmpl in (SF.conv2d_nhwc_cpu_template(), SF.matmul_mlt_cpu_template(256, 256, 256)):
        sk = SF.SketchFeatures(tmpl)
        for n in (1, 31, 65, 4097):
            e = torch.randint(1, 9, (n, tmpl.n_slots), dtype=torch.int32, device="cuda")
            r32 = sk.rows(e)
            r16 = sk.rows(e, half=True)
            assert r16.dtype == torch.float16 and torch.equal(r16, r32.clamp(-65504, 65504).half())
            off = torch.arange(0, (n + 1) * sk.n_rows, sk.n_rows, dtype=torch.int32, device="cuda")
            s32, s16 = torch.empty(n, device="cuda"), torch.empty(n, device="cuda")
            h = net.packed().h
            _lib.check(_lib.load().vaex_segsum_mlp_forward(h, _lib.ptr(r32), _lib.ptr(off), n, r32.shape[0], _lib.ptr(s32), 0, _lib.stream_ptr()))
            _lib.check(_lib.load().vaex_segsum_mlp_forward(h, _lib.ptr(r16), _lib.ptr(off), n, r16.shape[0], _lib.ptr(s16), _lib.F_FEATS_F16,
                                                           _lib.stream_ptr()))
            assert torch.equal(s32, s16)
