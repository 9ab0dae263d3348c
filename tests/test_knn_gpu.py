"""GPU: kernel (4) against the oracle / the reference's golden indices, through the C-ABI."""
import numpy as np
import pytest
import torch

from oracle import knn_oracle as KO
from plr2_b200 import synth
from plr2_b200.lib.knn import KNN_EXACT, KNN_FAST, KNearestNeighbor, knn_indices
from tests.util import knn_gap_ok, load_golden

pytestmark = pytest.mark.gpu


def run(ref, query, k=1, mode=KNN_EXACT, dist=False):
    r = torch.from_numpy(ref).cuda()[None].contiguous()
    q = torch.from_numpy(query).cuda()[None].contiguous()
    out = knn_indices(r, q, k, mode, return_dist=dist)
    torch.cuda.synchronize()
    if dist:
        return out[0][0].cpu().numpy(), out[1][0].cpu().numpy()
    return out[0].cpu().numpy()


@pytest.mark.parametrize("name", ["knn_d3_k1", "knn_d3_k4", "knn_d16_k2"])
def test_golden_reference_indices(name):
    g = load_golden(name)
    k = int(g["k"])
    idx = run(g["ref"], g["query"], k)
    _, best, second = KO.knn_oracle(g["ref"], g["query"], 1, contract=False, want_second=True)
    if k == 1:
        bad, decided, diff = knn_gap_ok(idx, g["idx"], best, second)
        assert bad == 0 and decided > 1000
        # exact ties planted in the fixture (duplicated refs, queries on refs) -> lowest index
        assert np.array_equal(idx[0, :8], g["idx"][0, :8])
    else:
        # generic kernel, fma arithmetic: compare with the fma oracle bit for bit
        assert np.array_equal(idx, KO.knn_oracle(g["ref"], g["query"], k, contract=True)[0])


@pytest.mark.parametrize("n_ref,n_query", [(1, 1), (1, 700), (15, 33), (16, 512), (17, 513), (500, 5000),
                                           (2047, 100), (2048, 1031), (2049, 777), (5000, 300),
                                           (4096, 1100), (4097, 3000), (9000, 2500)])
def test_exact_mode_bit_exact_vs_fma_oracle(n_ref, n_query):
    rng = np.random.default_rng(n_ref * 7 + n_query)
    ref = (rng.standard_normal((3, n_ref)) * 0.1 + 0.7).astype(np.float32)
    query = (rng.standard_normal((3, n_query)) * 0.1 + 0.7).astype(np.float32)
    idx, dist = run(ref, query, 1, KNN_EXACT, dist=True)
    oidx, odist = KO.knn_oracle(ref, query, 1, contract=True)
    assert np.array_equal(idx, oidx)
    assert np.array_equal(dist, odist)      # squared distances bit-identical to knn.cu arithmetic


def test_ties_on_lattice_lowest_index_wins():
    rng = np.random.default_rng(3)
    ref = rng.integers(-3, 4, (3, 777)).astype(np.float32)
    query = rng.integers(-3, 4, (3, 4099)).astype(np.float32)
    for mode in (KNN_EXACT, KNN_FAST):   # integers: both forms are exact -> identical indices
        assert np.array_equal(run(ref, query, 1, mode), KO.knn_oracle(ref, query, 1)[0])


def test_ties_on_lattice_multi_tile():
    """More references than one shared-memory tile holds (4096): rounds of 8/4/2/1 queries per lane re-stage
    the tiles, the winning chunk is re-read from global memory; lowest index still wins in both modes."""
    rng = np.random.default_rng(5)
    ref = rng.integers(-6, 7, (3, 9001)).astype(np.float32)
    query = rng.integers(-6, 7, (3, 2300)).astype(np.float32)
    for mode in (KNN_EXACT, KNN_FAST):
        assert np.array_equal(run(ref, query, 1, mode), KO.knn_oracle(ref, query, 1)[0])


def test_fast_mode_gap_rule_ycb_shape():
    fr = synth.make_frame(synth.YCB, 11, obj=12, with_image=False)
    pr, pt, _ = synth.make_predictions(fr, 12)
    # queries = transformed model points of the first 40 hypotheses (20 000 queries, 500 refs)
    qs = []
    for n in range(40):
        q = pr[n] / np.linalg.norm(pr[n])
        rot = synth.quat_to_rot_np(q).astype(np.float32)
        qs.append(fr["model_points"] @ rot.T + fr["points"][n] + pt[n])
    query = np.ascontiguousarray(np.concatenate(qs, 0).T.astype(np.float32))
    ref = np.ascontiguousarray(fr["target"].T)
    oidx, best, second = KO.knn_oracle(ref, query, 1, contract=True, want_second=True)
    for mode in (KNN_EXACT, KNN_FAST):
        bad, decided, diff = knn_gap_ok(run(ref, query, 1, mode), oidx, best, second)
        assert bad == 0 and decided > 0.9 * query.shape[1]
        if mode == KNN_EXACT:
            assert diff == 0


def test_nan_and_inf_follow_reference_kernel():
    ref = np.array([[0, 1, 2, 3], [0, 0, 0, 0], [0, 0, 0, 0]], np.float32)
    query = np.array([[np.nan, 2.2, np.inf], [0, 0, 0], [0, 0, 0]], np.float32)
    idx = run(ref, query)
    assert idx[0, 0] == 1          # all distances NaN: index stays at its initial 1 (knn.cu:127-128)
    assert idx[0, 1] == 3
    assert idx[0, 2] == 1          # all distances +inf: nothing is '<' the first
    ref2 = ref.copy(); ref2[0, 0] = np.nan          # dist[0] NaN -> every later '<' is false
    assert run(ref2, query[:, 1:2])[0, 0] == 1


def test_batched_and_dropin_signature():
    rng = np.random.default_rng(0)
    ref = rng.standard_normal((3, 3, 90)).astype(np.float32)
    query = rng.standard_normal((3, 3, 1000)).astype(np.float32)
    inds = KNearestNeighbor.apply(torch.from_numpy(ref), torch.from_numpy(query), 1)   # CPU in, like the reference
    assert inds.is_cuda and inds.dtype == torch.int64 and tuple(inds.shape) == (3, 1, 1000)
    inds_kw = KNearestNeighbor.apply(torch.from_numpy(ref).cuda(), torch.from_numpy(query).cuda(), k=1)
    assert torch.equal(inds, inds_kw)
    for b in range(3):
        assert np.array_equal(inds[b].cpu().numpy(), KO.knn_oracle(ref[b], query[b], 1)[0])
    # the reference's pybind entry, caller-allocated idx
    from plr2_b200.lib.knn import knn_pytorch
    out = torch.empty(3, 2, 1000, dtype=torch.int64, device="cuda")
    assert knn_pytorch.knn(torch.from_numpy(ref).cuda(), torch.from_numpy(query).cuda(), out) == 1
    assert np.array_equal(out[1].cpu().numpy(), KO.knn_oracle(ref[1], query[1], 2)[0])
    with pytest.raises(RuntimeError):
        knn_pytorch.knn(torch.from_numpy(ref).cuda().double(), torch.from_numpy(query).cuda(), out)


def test_full_size_properties_c2():
    """BASELINE config C2 size (500 refs x 500 000 queries): size-independent checks —
    (i) queries that ARE reference points return themselves (lowest duplicate), (ii) the returned
    distance is <= the distance to 64 random references, (iii) a 20 000-query slice equals the oracle."""
    rng = np.random.default_rng(42)
    n_ref, n_query = 500, 500_000
    ref = (rng.standard_normal((3, n_ref)) * 0.05).astype(np.float32)
    query = (rng.standard_normal((3, n_query)) * 0.05).astype(np.float32)
    self_ids = rng.integers(0, n_ref, 1000)
    query[:, :1000] = ref[:, self_ids]
    idx, dist = run(ref, query, 1, KNN_EXACT, dist=True)
    assert np.array_equal(idx[0, :1000], self_ids + 1) and np.all(dist[0, :1000] == 0)
    probe = rng.integers(0, n_ref, 64)
    d = ((query[:, :, None] - ref[:, None, probe]) ** 2).sum(0)
    assert np.all(dist[0][:, None] <= d * (1 + 1e-6) + 1e-12)
    sl = slice(300_000, 320_000)
    assert np.array_equal(idx[:, sl], KO.knn_oracle(ref, query[:, sl], 1)[0])
    fidx = run(ref, query, 1, KNN_FAST)
    assert (fidx != idx).mean() < 1e-3


# ---------------------------------------------------------------------------------------------
# The reference's OWN CUDA kernels (lib/knn/src/cuda/knn.cu, compiled unmodified for sm_100a into
# oracle/_ref/libknn_cuda_ref.so by oracle/Makefile) as the witness of bit-exactness.
needs_ref_cuda = pytest.mark.skipif(not KO.have_reference_cuda_knn(),
                                    reason="oracle/_ref/libknn_cuda_ref.so not built (make -C oracle where /root/reference exists)")


@needs_ref_cuda
@pytest.mark.parametrize("n_ref,n_query", [(500, 120_000), (2300, 60_000), (17, 4099), (1, 300)])
def test_exact_mode_bit_equal_to_reference_cuda_kernel(n_ref, n_query):
    """knn3_kernel<EXACT>: indices AND squared distances bit-equal to cuComputeDistanceGlobal + cuInsertionSort
    (knn.cu:36-176) run on the same B200, >= 1e5 queries, with planted exact ties (duplicated references, queries
    sitting on references, queries equidistant from two references)."""
    rng = np.random.default_rng(n_ref * 13 + n_query)
    ref = (rng.standard_normal((3, n_ref)) * 0.1 + 0.7).astype(np.float32)
    query = (rng.standard_normal((3, n_query)) * 0.1 + 0.7).astype(np.float32)
    if n_ref >= 17:
        ref[:, 11] = ref[:, 3]                                   # duplicated reference: index 4 (1-based) must win over 12
        ref[:, n_ref - 1] = ref[:, 5]
        query[:, :64] = ref[:, rng.integers(0, n_ref, 64)]       # queries on references (distance 0)
        mid = 0.5 * (ref[:, 7] + ref[:, 9])                      # equidistant in exact arithmetic, decided by rounding
        query[:, 64:96] = mid[:, None]
    r = torch.from_numpy(ref).cuda()[None].contiguous()
    q = torch.from_numpy(query).cuda()[None].contiguous()
    ridx, rdist = KO.knn_reference_cuda(r, q, 1, return_dist=True)
    idx, dist = knn_indices(r, q, 1, KNN_EXACT, return_dist=True)
    torch.cuda.synchronize()
    assert torch.equal(idx, ridx)
    assert torch.equal(dist[0], rdist)
    if n_ref >= 17:
        assert int(idx[0, 0, :64].max()) <= n_ref and not bool((idx[0, 0, :96] == 12).any())
    # and our C restatement of that arithmetic agrees with the real kernel too (pins oracle/knn_oracle.c, contract=1)
    sl = slice(0, min(n_query, 20_000))
    oidx, odist = KO.knn_oracle(ref, query[:, sl], 1, contract=True)
    assert np.array_equal(ridx[0, :, sl].cpu().numpy(), oidx) and np.array_equal(rdist[:, sl].cpu().numpy(), odist)


@needs_ref_cuda
def test_generic_k_and_dim_equal_reference_cuda_kernel():
    """k = 4 (dim 3) and k = 2 (dim 16) against the reference kernel: indices equal, batched."""
    rng = np.random.default_rng(17)
    for d, k, n_ref, n_query in ((3, 4, 300, 5000), (16, 2, 129, 3000)):
        ref = torch.from_numpy(rng.standard_normal((2, d, n_ref)).astype(np.float32)).cuda()
        query = torch.from_numpy(rng.standard_normal((2, d, n_query)).astype(np.float32)).cuda()
        assert torch.equal(knn_indices(ref, query, k, KNN_EXACT), KO.knn_reference_cuda(ref, query, k))


@needs_ref_cuda
def test_fast_mode_gap_rule_against_reference_cuda_kernel():
    """FAST ranking against the real reference kernel at C2 size: same neighbour wherever best and second-best
    squared distances (from the C oracle) differ by more than 1e-6."""
    rng = np.random.default_rng(99)
    ref = (rng.standard_normal((3, 500)) * 0.05 + np.array([[0.1], [-0.1], [0.9]])).astype(np.float32)
    query = (rng.standard_normal((3, 200_000)) * 0.05 + np.array([[0.1], [-0.1], [0.9]])).astype(np.float32)
    r, q = torch.from_numpy(ref).cuda()[None], torch.from_numpy(query).cuda()[None]
    ridx = KO.knn_reference_cuda(r, q, 1)[0].cpu().numpy()
    fidx = knn_indices(r, q, 1, KNN_FAST)[0].cpu().numpy()
    _, best, second = KO.knn_oracle(ref, query, 1, contract=True, want_second=True)
    bad, decided, diff = knn_gap_ok(fidx, ridx, best, second)
    assert bad == 0 and decided > 0.99 * query.shape[1]


def test_reference_wrapper_entry_with_non_contiguous_inputs():
    """knn_pytorch.knn on transposed (non-contiguous) ref AND query of equal byte size: the contiguous copies must
    both be alive when the kernel is enqueued (a freed ref copy could be handed to query.contiguous())."""
    from plr2_b200.lib.knn import knn_pytorch
    rng = np.random.default_rng(5)
    ref_t = torch.from_numpy(rng.standard_normal((2, 4000, 3)).astype(np.float32)).cuda()
    qry_t = torch.from_numpy(rng.standard_normal((2, 4000, 3)).astype(np.float32)).cuda()
    ref, qry = ref_t.transpose(1, 2), qry_t.transpose(1, 2)          # (2,3,4000) views, not contiguous
    assert not ref.is_contiguous() and not qry.is_contiguous()
    out = torch.empty(2, 1, 4000, dtype=torch.int64, device="cuda")
    for _ in range(3):
        assert knn_pytorch.knn(ref, qry, out) == 1
        for b in range(2):
            want = KO.knn_oracle(ref[b].contiguous().cpu().numpy(), qry[b].contiguous().cpu().numpy(), 1)[0]
            assert np.array_equal(out[b].cpu().numpy(), want)
