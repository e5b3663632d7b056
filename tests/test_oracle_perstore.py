"""Pins oracle/perstore_features.py -- the restated PerStoreFeatureExtractor (src/auto_scheduler/feature.cc:609-1188) -- to
the reference's own known-answer tests for it (tests/python/unittest/test_auto_scheduler_feature.py), with the lowered loop
nests those tests print in their comments written as templates.  CPU only."""
import math

import numpy as np

from oracle import perstore_features as PF
from vaex_b200 import sketch_features as SF


def fequal(a, b):
    return math.fabs(a - b) < 1e-5        # the reference uses 1e-6 on float32 values; log2 here is numpy's


def test_cpu_matmul_known_answers():
    """test_auto_scheduler_feature.py:31-95."""
    t = SF.matmul_cpu_template(512, 512, 512)
    rows, bufs = PF.evaluate(t.as_dict(), [32, 64, 512, 8, 16])
    names = PF.feature_names()
    assert rows.shape == (1, 164) and len(names) == 164
    fea = dict(zip(names, rows[0]))
    c_name = b_name = a_name = None
    for name in ["B0", "B1", "B2"]:
        if fequal(fea[name + ".acc_type.kReadWrite"], 1.0):
            c_name = name
        if fequal(fea[name + ".acc_type.kRead"], 1.0):
            if fequal(fea[name + ".stride"], 0.0):
                b_name = name
            else:
                a_name = name
    assert c_name and b_name and a_name
    assert fequal(fea[c_name + ".bytes"], math.log2(512 ** 3 * 4 + 1))
    assert fequal(fea[b_name + ".unique_bytes"], math.log2(512 ** 2 * 4 + 1))
    assert fequal(fea[c_name + ".reuse_dis_iter"], math.log2(8 * 16 + 1))
    assert fequal(fea[c_name + ".reuse_dis_bytes"], math.log2((8 * 16 + 8 + 16) * 4 + 1))
    assert fequal(fea[c_name + ".reuse_ct"], math.log2(512 + 1))
    assert fequal(fea["unroll_num"], math.log2(1 + 1))
    assert fequal(fea["vec_num"], math.log2(1 + 1))
    assert fequal(fea["parallel_num"], math.log2(2 + 1))
    assert fequal(fea["parallel_prod"], math.log2((512 * 512 / 16 / 8) + 1))
    # not asserted by the reference but implied by its loop nest
    assert fequal(fea["float_mul"], math.log2(512 ** 3 + 1)) and fequal(fea["num_loops"], math.log2(5 + 1))
    assert fea["vec_type.kPosMixed"] == 1.0 and fea["is_gpu"] == 0.0


def fusion_template(N, M):
    """test_auto_scheduler_feature.py:97-124: B = A, C = B with B computed at C's inner loop (its own loops have extent 1)."""
    t = SF.SketchTemplate(2)
    A, B, Cb = t.buffer("A", (N, M)), t.buffer("B", (N, M)), t.buffer("C", (N, M))
    one = t.const(1)
    with t.loop("i", t.slot(0)) as i:
        with t.loop("j", t.slot(1)) as j:
            with t.realize(B, [one, one]):
                with t.loop("ii", one) as ii:
                    with t.loop("jj", one) as jj:
                        ix = [SF.add(SF.V(i), SF.V(ii)), SF.add(SF.V(j), SF.V(jj))]
                        t.store(B, ix, reads=[(A, ix)])
            t.store(Cb, [SF.V(i), SF.V(j)], reads=[(B, [SF.V(i), SF.V(j)])])
    return t


def test_cpu_fusion_known_answers():
    """test_auto_scheduler_feature.py:97-137."""
    rows, _ = PF.evaluate(fusion_template(64, 32).as_dict(), [64, 32])
    names = PF.feature_names()
    found = False
    for row in rows:
        for i, (name, value) in enumerate(zip(names, row)):
            if "reuse_type.kSerialMultipleReadWrite" in name and value > 0.5:
                assert fequal(row[i + 2], 1.0)
                assert fequal(row[i + 3], math.log2(8 + 1))
                assert fequal(row[i + 4], math.log2(64 * 32 + 1))
                found = True
    assert found


def matmul_gpu_template():
    """The lowered loop nest printed at test_auto_scheduler_feature.py:173-195 (512^3 matmul, multi-level tiling with
    shared-memory cache reads); extents fixed, indices written out from the split factors of the record above it."""
    t = SF.SketchTemplate(1)
    c = t.const
    A, B = t.buffer("A", (512, 512)), t.buffer("B", (512, 512))
    Cl, As, Bs, Cb = t.buffer("C.local", (512, 512)), t.buffer("A.shared", (512, 512)), t.buffer("B.shared", (512, 512)), t.buffer("C", (512, 512))
    with t.thread("blockIdx.x", c(8)) as bx:
        with t.thread("vthread", c(4)) as vt:
            with t.thread("threadIdx.x", c(16)) as tx:
                # i = i.0 (1) i.1 (1) i.2 (16) i.3 (32);  j = j.0 (8 -> blockIdx) j.1 (4 -> vthread) j.2 (1) j.3 (16)
                ci = lambda i3: SF.add(SF.mul(SF.V(tx), SF.C(c(32))), SF.V(i3))
                cj = lambda j3: SF.add(SF.add(SF.mul(SF.V(bx), SF.C(c(64))), SF.mul(SF.V(vt), SF.C(c(16)))), SF.V(j3))
                with t.realize(Cl, [c(32), c(16)]):
                    with t.pragma_unroll(c(1024)):
                        with t.loop("i_c.3.init", c(32)) as ii:
                            with t.loop("j_c.4.init", c(16)) as ji:
                                t.store(Cl, [ci(ii), cj(ji)])
                        with t.loop("k.0", c(256)) as k0:
                            with t.realize(Bs, [c(2), c(64)]):
                                with t.loop("ax0@ax1@.0", c(8)) as f0:
                                    with t.thread("threadIdx.x", c(16)) as f1:
                                        f = SF.add(SF.mul(SF.V(f0), SF.C(c(16))), SF.V(f1))
                                        ix = [SF.add(SF.mul(SF.V(k0), SF.C(c(2))), SF.floordiv(f, SF.C(c(64)))),
                                              SF.add(SF.mul(SF.V(bx), SF.C(c(64))), SF.floormod(f, SF.C(c(64))))]
                                        t.store(Bs, ix, reads=[(B, ix)])
                            with t.realize(As, [c(512), c(2)]):
                                with t.loop("ax0@ax1@.0", c(64)) as g0:
                                    with t.thread("threadIdx.x", c(16)) as g1:
                                        g = SF.add(SF.mul(SF.V(g0), SF.C(c(16))), SF.V(g1))
                                        ix = [SF.floordiv(g, SF.C(c(2))), SF.add(SF.mul(SF.V(k0), SF.C(c(2))), SF.floormod(g, SF.C(c(2))))]
                                        t.store(As, ix, reads=[(A, ix)])
                            with t.loop("i_c.3", c(32)) as i3:
                                with t.loop("k.2", c(2)) as k2:
                                    with t.loop("j_c.4", c(16)) as j4:
                                        kk = SF.add(SF.mul(SF.V(k0), SF.C(c(2))), SF.V(k2))
                                        t.store(Cl, [ci(i3), cj(j4)], reads=[(Cl, [ci(i3), cj(j4)]), (As, [ci(i3), kk]), (Bs, [kk, cj(j4)])],
                                                ops={"float_addsub": 1, "float_mul": 1})
                with t.loop("i.3", c(32)) as o3:
                    with t.loop("j.3", c(16)) as p3:
                        t.store(Cb, [ci(o3), cj(p3)], reads=[(Cl, [ci(o3), cj(p3)])])
    return t


def test_gpu_matmul_known_answers():
    """test_auto_scheduler_feature.py:139-204."""
    rows, bufs = PF.evaluate(matmul_gpu_template().as_dict(), [1])
    names = PF.feature_names()
    assert rows.shape == (4, 164)
    for row in rows:
        fea = dict(zip(names, row))
        assert fequal(fea["blockIdx_x_len"], math.log2(8 + 1)) and fequal(fea["vthread_len"], math.log2(4 + 1))
        assert fequal(fea["threadIdx_x_len"], math.log2(16 + 1)) and fequal(fea["threadIdx_y_len"], math.log2(1 + 1))
        assert fequal(fea["blockIdx_z_len"], math.log2(1 + 1)) and fea["is_gpu"] == 1.0
    local = dict(zip(names, rows[bufs.index(2)]))
    assert fequal(local["auto_unroll_max_step"], math.log2(1024 + 1))
    assert fequal(local["alloc_size"], math.log2(32 * 16 * 4 + 1))


def test_trivial_loops_vanish():
    """A loop of extent 1 is not in the lowered nest (ScheduleOps, debug_keep_trivial_loop = false): the rows of the
    matmul template with i.1 = 1 equal those of the same nest written without that loop."""
    t = SF.matmul_cpu_template(64, 64, 64)
    rows, _ = PF.evaluate(t.as_dict(), [64, 8, 64, 8, 1])
    fea = dict(zip(PF.feature_names(), rows[0]))
    assert fequal(fea["num_loops"], math.log2(4 + 1)) and fequal(fea["outer_prod"], math.log2(64 * 8 * 64 * 8 + 1))
    assert np.isfinite(rows).all()
