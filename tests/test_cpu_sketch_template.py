"""Host side of n4 (no GPU): the template builder / compiler of vae-extent-search_b200/sketch_features.py -- program layout,
term recognition, limits -- and template-level properties checked through the oracle."""
import numpy as np
import pytest

from oracle import perstore_features as PF
from vaex_b200 import sketch_features as SF


def test_program_layout_and_limits():
    t = SF.conv2d_nhwc_cpu_template()
    prog = t.compile()
    assert prog.dtype == np.int32 and prog[0] == SF.MAGIC and prog[1] == 23 and prog[5] == 3 == t.n_rows
    assert prog[2] == len(t.scalars) <= SF.MAX_SCALARS and prog[3] == 6 and len(prog) <= SF.MAX_PROG
    offs = prog[7:12]
    assert list(offs) == sorted(offs) and offs[-1] < len(prog)
    n_instr = prog[6]
    instr = prog[offs[2]:offs[2] + 4 * n_instr].reshape(-1, 4)
    stores = instr[instr[:, 0] == SF.OP_STORE]
    assert len(stores) == 4 and stores[:, 2].tolist() == [1, 0, 1, 1]       # init, update of Conv2dOutput: only the update keeps a row
    assert (instr[:, 0] == SF.OP_FOR).sum() == (instr[:, 0] == SF.OP_FOR_END).sum() == 36
    assert instr[instr[:, 0] == SF.OP_FOR][:, 1].max() < SF.MAX_DEPTH        # loop ids are nesting depths
    with pytest.raises(ValueError):
        SF.SketchTemplate(0)
    with pytest.raises(ValueError):
        SF.SketchTemplate(4).scalar(1, (4,))
    with pytest.raises(ValueError):
        SF.SketchTemplate(4).buffer("x", (1, 2, 3, 4, 5))
    deep = SF.SketchTemplate(1)
    b = deep.buffer("b", (4,))
    with deep.loops(*[("l%d" % i, deep.slot(0)) for i in range(SF.MAX_DEPTH + 1)]) as ls:
        deep.store(b, [SF.V(ls[-1])])
    with pytest.raises(ValueError, match="nested deeper"):
        deep.compile()


def test_term_recognition():
    t = SF.SketchTemplate(2)
    with t.loop("a", t.slot(0)) as a, t.loop("b", t.slot(1)) as b:
        c4, c3 = t.const(4), t.const(3)
        canon = SF.lin([(SF.floormod(SF.floordiv(SF.V(a), SF.C(c4)), SF.C(c3)), c4), (SF.V(b), c3), (SF.V(a), None)])
        terms = SF._terms(canon)
        assert [k for k, *_ in terms] == [SF.TERM_DIVMOD, SF.TERM_PLAIN, SF.TERM_PLAIN] and terms[0][2:] == (c4, c4, c3)
        assert SF._terms(SF.add(SF.V(a), SF.C(c4)))[1][0] == SF.TERM_CONST
        assert SF._terms(SF.floordiv(SF.add(SF.mul(SF.V(a), SF.C(c4)), SF.V(b)), SF.C(c3))) is None      # floordiv of a sum: token form
        assert SF._terms(SF.mul(SF.V(a), SF.V(b))) is None
        assert SF._terms(SF.add(SF.V(a), SF.add(SF.V(b), SF.C(c4)))) is None                              # right-nested sum: token form


def test_derived_outer_levels_equal_explicit_ones():
    """conv2d template: the outermost tile level derived as axis length // product(inner levels) gives the rows of the
    explicit 23-slot form, for legal tilings."""
    full, derived = SF.conv2d_nhwc_cpu_template(), SF.conv2d_nhwc_cpu_template(derive_outer=True)
    rng = np.random.default_rng(0)
    length = {"n": 1, "y": 28, "x": 28, "f": 128, "ry": 3, "rx": 3, "rc": 64}
    for _ in range(12):
        e = np.ones(23, np.int64)
        for key, E in length.items():
            sl, rest = SF.CONV2D_SLOTS[key], E
            for j in sl[:0:-1]:
                divs = [d for d in range(1, rest + 1) if rest % d == 0]
                e[j] = rng.choice(divs)
                rest //= e[j]
            e[sl[0]] = rest
        e[22] = rng.choice([0, 16, 64, 512])
        a, _ = PF.evaluate(full.as_dict(), e.tolist())
        junk = e.copy()
        junk[[0, 4, 8, 12, 16, 18, 20]] = 777                     # the derived form never reads the outer slots
        b, _ = PF.evaluate(derived.as_dict(), junk.tolist())
        assert np.array_equal(a, b)
    se, sl_ = SF.conv2d_slot_tables()
    assert se[[5, 6, 7]].tolist() == [28, 28, 28] and se[[0, 4, 8, 12, 16, 18, 20, 22]].tolist() == [0] * 8
    assert sl_.nonzero()[0].tolist() == [3, 7, 11, 15, 17, 19, 21]


def test_rows_are_invariant_to_what_the_extractor_cannot_see():
    """Properties of the extractor on the matmul template: the total work (float_mul = N * M * K) does not depend on the
    tiling; swapping in a trivial tile leaves the row of the same nest written without that loop (checked in
    test_oracle_perstore); every value is finite and the one-hot groups sum to one."""
    t = SF.matmul_cpu_template(256, 256, 256)
    names = PF.feature_names()
    for e in ([16, 32, 256, 8, 16], [256, 256, 256, 1, 1], [1, 1, 256, 256, 256], [4, 2, 256, 128, 64]):
        rows, _ = PF.evaluate(t.as_dict(), e)
        fea = dict(zip(names, rows[0]))
        assert np.isfinite(rows).all()
        assert abs(fea["float_mul"] - np.log2(256.0 ** 3 + 1)) < 1e-5
        for grp in ("vec_type.", "unroll_type.", "parallel_type.", "B0.acc_type.", "B0.reuse_type."):
            assert sum(v for k, v in fea.items() if k.startswith(grp)) == 1.0


def test_conv2d_template_reproduces_the_readme_program():
    """The one lowered-program listing the reference ships for this sketch (README.md:86-108, kept as a fixture in
    oracle/gen_golden.py): with its extents the template's non-trivial loops are that listing, loop for loop -- annotation,
    extent, nesting and the statement each nest ends in.  (PrintState does not list the reduction's init nest.)"""
    import re
    from oracle.gen_golden import README_STATE
    want = []
    for line in README_STATE.splitlines()[1:]:
        ind = (len(line) - len(line.lstrip())) // 2
        m = re.match(r"\s*(parallel|vectorize|unroll|for) \S+ \(.*?(\d+)\)\s*$", line)
        if m:
            want.append((ind, m.group(1), int(m.group(2))))
        elif "auto_unroll" in line:
            want.append((ind, "pragma", int(line.rsplit(":", 1)[1])))
        elif line.strip().endswith("= ..."):
            want.append((ind, "store", line.split("=")[0].strip()))
    t = SF.conv2d_nhwc_cpu_template()
    S = PF.scalar_values(t.as_dict(), SF.CONV2D_README_EXTENTS)
    got = []

    def walk(nodes, ind):
        for n in nodes:
            k = n["kind"]
            if k in SF.KINDS:
                if S[n["extent"]] == 1 or n["name"].endswith(".init"):
                    walk(n["body"], ind)                      # trivial loops are not printed
                    continue
                got.append((ind, k, S[n["extent"]]))
                walk(n["body"], ind + 1)
            elif k == "pragma_unroll":
                got.append((ind, "pragma", S[n["value"]]))
                walk(n["body"], ind)
            elif k == "realize":
                walk(n["body"], ind)
            elif k == "store":
                if n.get("reads"):
                    got.append((ind, "store", t.buffers[n["buffer"]]["name"]))
    walk(t.body, 0)
    assert got == want, (got, want)


def test_conv2d_template_matches_task_1264s_flop_count():
    """README.md:63-66: task 1264 (the conv2d + bias + ReLU workload the template states) has flop_ct 115806208.  The
    template's statements -- ops per BufferStore times the trip count of its nest, the way the extractor counts them --
    give exactly that (the if_then_else of the padding stage is a call, which TVM's FLOP estimator does not count)."""
    t = SF.conv2d_nhwc_cpu_template()
    fs = PF.feature_sets(t.as_dict(), SF.CONV2D_README_EXTENTS)
    flops = sum(float(f[k]) for f in fs.values() for k in ("float_mad", "float_addsub", "float_mul", "float_divmod", "float_cmp"))
    assert flops == 115806208.0
    assert sorted(t.buffers[b]["name"] for b in fs) == ["Conv2dOutput", "PaddedInput", "T_relu"]
    pad = fs[[b for b in fs if t.buffers[b]["name"] == "PaddedInput"][0]]
    assert float(pad["int_cmp"]) == 4 * 218880 and float(pad["bool_op"]) == 3 * 218880 and float(pad["float_math_func"]) == 218880


def test_matmul_multi_level_tiling_template():
    """Work invariants of the cache-write matmul sketch: 2 rows (C.local, C); M * N * K multiplies and M * N stores whatever the
    tiling; the derived-outer form equals the explicit one on legal tilings."""
    full, derived = SF.matmul_mlt_cpu_template(256, 128, 64), SF.matmul_mlt_cpu_template(256, 128, 64, derive_outer=True)
    names = PF.feature_names()
    for e in ([4, 4, 4, 4, 2, 2, 4, 8, 8, 8, 64], [256, 1, 1, 1, 1, 1, 1, 128, 1, 64, 0], [1, 1, 1, 256, 128, 1, 1, 1, 64, 1, 512]):
        rows, bufs = PF.evaluate(full.as_dict(), e)
        assert [full.buffers[b]["name"] for b in bufs] == ["C.local", "C"]
        local, out = dict(zip(names, rows[0])), dict(zip(names, rows[1]))
        assert abs(local["float_mul"] - np.log2(256 * 128 * 64 + 1.0)) < 1e-5 and abs(out["outer_prod"] - np.log2(256 * 128 + 1.0)) < 1e-5
        junk = list(e)
        junk[0] = junk[4] = junk[8] = 999
        rows_d, _ = PF.evaluate(derived.as_dict(), junk)
        assert np.array_equal(rows, rows_d)


# The transform steps of a measured llvm program of the same workload function (conv2d NHWC + bias + ReLU, hash 12b88bed...)
# shipped with the reference: tutorials/auto_scheduler/ci_logs/resnet-50-NHWC-B1-llvm.json, line 14.
CI_LOG_STEPS = [["CI", 5], ["SP", 3, 0, 1, [1, 1, 1], 1], ["SP", 3, 4, 112, [1, 1, 4], 1], ["SP", 3, 8, 112, [4, 2, 1], 1],
                ["SP", 3, 12, 64, [1, 1, 16], 1], ["SP", 3, 16, 7, [7], 1], ["SP", 3, 18, 7, [7], 1], ["SP", 3, 20, 3, [3], 1],
                ["RE", 3, [0, 4, 8, 12, 1, 5, 9, 13, 16, 18, 20, 2, 6, 10, 14, 17, 19, 21, 3, 7, 11, 15]],
                ["FSP", 6, 0, 1, 2], ["FSP", 6, 3, 2, 2], ["FSP", 6, 6, 3, 2], ["FSP", 6, 9, 4, 2],
                ["RE", 6, [0, 3, 6, 9, 1, 4, 7, 10, 2, 5, 8, 11]], ["CA", 3, 6, 7], ["CA", 1, 6, 3], ["FU", 6, [0, 1, 2]],
                ["AN", 6, 0, 3], ["PR", 3, 0, "auto_unroll_max_step$64"], ["AN", 1, 3, 2], ["AN", 3, 21, 2], ["AN", 6, 9, 2]]


def test_conv2d_template_has_the_structure_of_a_measured_ansor_record():
    """The sketch the template states is the one Ansor's CPU policy produced for this workload in the reference's own logs:
    the order of the 22 tile loops of Conv2dOutput (the record's reorder step: S S R S R S with ry, rx, rc inside each reduction
    level), the consumer tiled to follow the two outer levels, Conv2dOutput attached under the consumer's second-level channel
    loop, the three outer spatial tiles fused and parallelised, the innermost channel tiles vectorised, the unroll pragma on
    Conv2dOutput.  (Where the padding stage is attached and how it is annotated varies from record to record.)"""
    steps = CI_LOG_STEPS
    # iterators of stage 3 after its seven splits: n 0-3, yy 4-7, xx 8-11, ff 12-15, ry 16-17, rx 18-19, rc 20-21
    name = {}
    for base, ax, levels in ((0, "nn", 4), (4, "yy", 4), (8, "xx", 4), (12, "ff", 4), (16, "ry", 2), (18, "rx", 2), (20, "rc", 2)):
        for l in range(levels):
            name[base + l] = "%s.%d" % (ax, l)
    conv_order = [name[i] for i in [s for s in steps if s[0] == "RE" and s[1] == 3][0][2]]
    cons = {}
    for base, ax in ((0, "ax0"), (3, "ax1"), (6, "ax2"), (9, "ax3")):
        for l in range(3):
            cons[base + l] = "%s.%d" % (ax, l)
    cons_order = [cons[i] for i in [s for s in steps if s[0] == "RE" and s[1] == 6][0][2]]
    assert cons_order[:8] == ["ax0.0", "ax1.0", "ax2.0", "ax3.0", "ax0.1", "ax1.1", "ax2.1", "ax3.1"]
    assert ["CA", 3, 6, 7] in steps and cons_order[7] == "ax3.1"            # Conv2dOutput computed at the consumer's ax3.1
    assert ["FU", 6, [0, 1, 2]] in steps and ["AN", 6, 0, 3] in steps       # n.0, y.0, x.0 fused, parallel
    t = SF.conv2d_nhwc_cpu_template()
    path = {}                                                                # store buffer name (update store) -> enclosing loops

    def walk(nodes, stack):
        for n in nodes:
            if n["kind"] in SF.KINDS:
                walk(n["body"], stack + [(n["name"], n["kind"])])
            elif n["kind"] in ("realize", "pragma_unroll"):
                walk(n["body"], stack + ([("pragma", "pragma")] if n["kind"] == "pragma_unroll" else []))
            elif n["kind"] == "store" and n.get("reads"):
                path[t.buffers[n["buffer"]]["name"]] = stack
    walk(t.body, [])
    conv = path["Conv2dOutput"]
    names = [nm for nm, _ in conv]
    # consumer's outer loops (fused n.0 y.0 x.0, then ax3.0, ax0.1 .. ax3.1), the pragma, then Conv2dOutput's own 14 inner loops
    assert names[:6] == ["ax0.0@ax1.0@ax2.0@", "ax3.0", "ax0.1", "ax1.1", "ax2.1", "ax3.1"] and conv[0][1] == "parallel"
    assert names[6] == "pragma"
    assert names[7:] == conv_order[8:]                                       # ry.0 rx.0 rc.0 | nn.2 .. ff.2 | ry.1 rx.1 rc.1 | nn.3 .. ff.3
    assert conv[-1] == ("ff.3", "vectorize") and ["AN", 3, 21, 2] in steps and conv_order[21] == "ff.3"    # iterator 21 after the reorder
    relu = path["T_relu"]
    assert [nm for nm, _ in relu] == ["ax0.0@ax1.0@ax2.0@", "ax3.0", "ax0.1", "ax1.1", "ax2.1", "ax3.1", "ax0.2", "ax1.2", "ax2.2", "ax3.2"]
    assert relu[-1][1] == "vectorize" and ["AN", 6, 9, 2] in steps          # after the fuse, iterator 9 of the consumer is ax3.2
    assert [s for s in steps if s[0] == "PR"] == [["PR", 3, 0, "auto_unroll_max_step$64"]]


def test_template_and_extents_from_the_measured_record():
    """conv2d_template_from_record on the reference's ci_logs record (224 x 224 x 3 -> 112 x 112 x 64, 7 x 7, stride 2): the
    shapes and the padding (3 before, 2 after) from the workload key, the tile extents from the SplitSteps (each axis multiplies
    back to its length), the padding stage attached one loop further in and vectorised as that record has it; the rows the
    extractor then gives carry the workload's operation count."""
    key = '["12b88bedece6984af589a28b43e0f3c4", 1, 224, 224, 3, 7, 7, 3, 64, 1, 1, 1, 64, 1, 112, 112, 64]'
    rec = {"i": [[key, "llvm -keys=cpu -link-params=0 -mcpu=core-avx2"], [[], CI_LOG_STEPS]]}
    t, ext = SF.conv2d_template_from_record(rec)
    assert ext == [1, 1, 1, 1, 28, 1, 1, 4, 14, 4, 2, 1, 4, 1, 1, 16, 1, 7, 1, 7, 1, 3, 64]
    for key, E in (("n", 1), ("y", 112), ("x", 112), ("f", 64), ("ry", 7), ("rx", 7), ("rc", 3)):
        assert int(np.prod([ext[j] for j in SF.CONV2D_SLOTS[key]])) == E
    assert t.buffers[3]["name"] == "PaddedInput" and t.buffers[3]["shape"] == [1, 229, 229, 3]
    fs = PF.feature_sets(t.as_dict(), ext)
    flops = sum(float(f[k]) for f in fs.values() for k in ("float_addsub", "float_mul", "float_cmp"))
    assert flops == 2.0 * 112 * 112 * 64 * 7 * 7 * 3 + 2.0 * 112 * 112 * 64
    pad = fs[3]
    assert float(pad["vec_len"]) == 3.0 and float(pad["num_loops"]) == 5.0      # fused, ax3.0, i1, i2, i3 (vectorised); i0 is trivial
    bad = {"i": [[key, "llvm"], [[], [s for s in CI_LOG_STEPS if s[0] != "RE"]]]}
    with pytest.raises(ValueError):
        SF.conv2d_template_from_record(bad)


def test_ranked_candidates_go_back_out_as_measure_records():
    """conv2d_sketch_slots: the template's extent vector written back into the record's SplitSteps / pragma -- the record's
    own extents reproduce its steps, a different legal tiling changes exactly the lengths, an illegal one is refused."""
    import json
    key = '["12b88bedece6984af589a28b43e0f3c4", 1, 224, 224, 3, 7, 7, 3, 64, 1, 1, 1, 64, 1, 112, 112, 64]'
    rec = {"i": [[key, "llvm -keys=cpu -link-params=0 -mcpu=core-avx2"], [[], CI_LOG_STEPS]], "r": [[0.00025], 0, 0.3, 1606961659], "v": "v0.5"}
    slots = SF.conv2d_sketch_slots(rec)
    t, ext = SF.conv2d_template_from_record(rec)
    back = json.loads(slots.records_from_extents(np.array([ext]))[0])
    assert back["i"][1][1] == CI_LOG_STEPS and back["i"][0] == rec["i"][0]
    other = list(ext)
    other[5:8] = [2, 2, 7]                         # y: 112 = 4 * 2 * 2 * 7
    other[22] = 512
    steps = json.loads(slots.records_from_extents(np.array([other]))[0])["i"][1][1]
    assert steps[2] == ["SP", 3, 4, 112, [2, 2, 7], 1] and steps[18] == ["PR", 3, 0, "auto_unroll_max_step$512"]
    assert [s for k, s in enumerate(steps) if k not in (2, 18)] == [s for k, s in enumerate(CI_LOG_STEPS) if k not in (2, 18)]
    other[5:8] = [3, 2, 7]
    with pytest.raises(ValueError):
        slots.records_from_extents(np.array([other]))
    se, sl = slots.legalise_tables()[:2]
    want_se, want_sl = SF.conv2d_slot_tables(1, 224, 224, 3, 64, 7, 7, 2, 3, 2)
    assert np.array_equal(se, want_se) and np.array_equal(sl, want_sl)
