"""gcr_score_models_static (include/gcr_b200.h): the allocation-free variant of the batched FP64 model scoring -- scratch in a file-scope
`static __device__` array, no engine handle, device pointers in and out -- against the existing entry point gcr_score_models, through the
C ABI, on two calls in a row (the shared array keeps the first call's contents: the second call, on another scene with other sizes, must
not see them).  Both run the same FP64 reference-order kernels, so scores are compared to 1e-12 relative (far inside the 1e-5 bar of the
parity tests) and counts / packed masks bit for bit.  Collected last on purpose."""
import ctypes as C
import math

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _static_call(lib, torch, d_scale, d_orient, d_models, thr, flags=0, want_masks=True, stream=0):
    ns, no, n = d_scale.shape[0], d_orient.shape[0], d_models.shape[0]
    W = (ns + 31) // 32 + (no + 31) // 32
    score = torch.full((n,), float("nan"), dtype=torch.float64, device="cuda")
    cnt = torch.full((n, 2), -1, dtype=torch.int32, device="cuda")
    masks = torch.full((n, W), -1, dtype=torch.int32, device="cuda") if want_masks else None
    err = C.create_string_buffer(512)
    thr_c = (C.c_double * 2)(float(thr[0]), float(thr[1]))
    rc = lib.gcr_score_models_static(0, d_scale.data_ptr(), ns, d_orient.data_ptr(), no, d_models.data_ptr(), n, thr_c, flags,
                                     score.data_ptr(), cnt.data_ptr(), masks.data_ptr() if want_masks else None, C.c_void_p(stream), err, 512)
    return rc, err.value.decode(), score, cnt, masks


def test_static_scratch_scoring_equals_engine_scoring_twice_in_a_row():
    import torch
    import gcr_capi
    import gcr_scenes as gs
    lib = gcr_capi.load_library()
    eng = gcr_capi.Engine(0)
    rng = np.random.default_rng(11)
    cases = [(gs.make_scene(3000, 900, 0.3, 501), 300, 0), (gs.make_scene(700, 1300, 0.5, 502, thr_scale=0.1, thr_orient=math.radians(1.5)), 64, gcr_capi.SCORE_TRUNCATED_SQ)]
    for k, (sc, n, flags) in enumerate(cases):
        # models around the planted plane (plenty of inliers) plus a few far-off ones
        models = np.tile(sc.model, (n, 1)) * (1.0 + 0.02 * rng.standard_normal((n, 4)))
        models[::7, :2] *= -3.0
        eng.set_features(sc.scale, sc.orient)
        want_score, want_cnt, want_words = eng.score_models(models, sc.thr, truncated_sq=bool(flags), masks=True)
        d_scale, d_orient = torch.from_numpy(sc.scale).cuda(), torch.from_numpy(sc.orient).cuda()
        d_models = torch.from_numpy(models).cuda()
        torch.cuda.synchronize()
        rc, msg, score, cnt, masks = _static_call(lib, torch, d_scale, d_orient, d_models, sc.thr, flags)
        assert rc == 0, msg
        torch.cuda.synchronize()
        got_score, got_cnt = score.cpu().numpy(), cnt.cpu().numpy().view(np.uint32)
        got_words = masks.cpu().numpy().view(np.uint32)
        assert np.array_equal(got_cnt, want_cnt), k
        assert np.array_equal(got_words, want_words), k
        assert np.allclose(got_score, want_score, rtol=1e-12, atol=0.0), k
        assert want_cnt.sum() > 0  # the comparison is not about empty results
        # outputs are optional
        rc, msg, score2, _, _ = _static_call(lib, torch, d_scale, d_orient, d_models, sc.thr, flags, want_masks=False)
        assert rc == 0, msg
        torch.cuda.synchronize()
        assert np.array_equal(score2.cpu().numpy(), got_score)
    eng.close()


def test_static_scratch_rejects_sizes_beyond_its_maximum_and_allocates_nothing():
    import torch
    import gcr_capi
    import gcr_scenes as gs
    lib = gcr_capi.load_library()
    sc = gs.make_scene(2000, 500, 0.3, 503)
    d_scale, d_orient = torch.from_numpy(sc.scale).cuda(), torch.from_numpy(sc.orient).cuda()
    d_models = torch.from_numpy(np.tile(sc.model, (1025, 1))).cuda()
    rc, msg, *_ = _static_call(lib, torch, d_scale, d_orient, d_models, sc.thr)
    assert rc == gcr_capi.ERR_INVALID and "1024" in msg
    # 70 000 features: 2 189 mask words per model -> at most 110 models fit the 1 MiB array
    big = gs.make_scene(50_000, 20_000, 0.1, 504)
    db_scale, db_orient = torch.from_numpy(big.scale).cuda(), torch.from_numpy(big.orient).cuda()
    rc, msg, *_ = _static_call(lib, torch, db_scale, db_orient, d_models[:200], big.thr)
    assert rc == gcr_capi.ERR_INVALID and "exceeds the static scratch" in msg
    rc, msg, score, cnt, _ = _static_call(lib, torch, db_scale, db_orient, torch.from_numpy(np.tile(big.model, (110, 1))).cuda(), big.thr)
    assert rc == 0, msg
    torch.cuda.synchronize()
    assert (cnt.cpu().numpy()[:, 0] > 1000).all() and np.isfinite(score.cpu().numpy()).all()
    # no device memory is allocated by a call: free memory as the driver reports it is unchanged across calls
    out_s = torch.empty(64, dtype=torch.float64, device="cuda")
    out_c = torch.empty((64, 2), dtype=torch.int32, device="cuda")
    thr_c = (C.c_double * 2)(*sc.thr)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for _ in range(5):
        rc = lib.gcr_score_models_static(0, d_scale.data_ptr(), 2000, d_orient.data_ptr(), 500, d_models.data_ptr(), 64, thr_c, 0,
                                         out_s.data_ptr(), out_c.data_ptr(), None, None, None, 0)
        assert rc == 0
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0
    rc = lib.gcr_score_models_static(0, None, 2000, d_orient.data_ptr(), 500, d_models.data_ptr(), 64, thr_c, 0, None, None, None, None, None, 0)
    assert rc == gcr_capi.ERR_INVALID
