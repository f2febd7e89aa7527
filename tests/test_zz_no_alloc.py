"""GPU (-m gpu): nldsc_ld_snp_stats_static, the allocation-free entry point (static __device__ scratch), called through the C ABI
and checked against the existing entry points on the same rows: `maf` / `residuals_std` of nldsc_ld_calculate_packed and the
genotype counts of nldsc_ld_debug_unpack.  Bar: bit-equal (MAF, the counts and the closed-form residual std come from the same
integer counts through the same stats kernel).  Several calls in a row share the scratch array: a smaller, different case after a
larger one, then the first case again, prove that a call clears what it accumulates into.

Kept in a file of its own that pytest collects LAST."""
import numpy as np
import pytest
import torch

from nldsc_b200 import build, capi, synth

pytestmark = pytest.mark.gpu

MAF, STD = 1e-4, 1e-5


@pytest.fixture(scope="module", autouse=True)
def built():
    build.build_module()
    assert capi.device_count() >= 1


def _case(n_org, n_snp, missing, seed):
    spec = synth.SynthSpec(n_org=n_org, n_snp=n_snp, missing=missing, seed=seed)
    packed = synth.generate_packed(spec)
    cm, _ = synth.genetic_map(spec)
    packed[3] = 0                                   # a monomorphic SNP: MAF 0 -> dropped, residual std NaN
    want = capi.calculate_packed(packed, n_snp, n_org, cm, 1.0, MAF, STD, 1.0 / n_snp)
    _, counts = capi.debug_unpack(packed, n_org)
    return {"packed": packed, "n_org": n_org, "n_snp": n_snp, "maf": want["maf"], "rstd": want["residuals_std"],
            "counts": counts}


def _static(g, with_counts=True):
    """One call of the new entry point: device tensors in and out (torch owns them; the entry point allocates nothing)."""
    n = g["n_snp"]
    d_packed = torch.from_numpy(g["packed"]).cuda()
    d_counts = torch.full((n, 3), -7, dtype=torch.int32, device="cuda")
    d_maf = torch.full((n,), -7.0, dtype=torch.float64, device="cuda")
    d_rstd = torch.full((n,), -7.0, dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    capi.snp_stats_static(d_packed.data_ptr(), n, g["n_org"], MAF, STD, d_counts.data_ptr() if with_counts else 0,
                          d_maf.data_ptr(), d_rstd.data_ptr(), device=0, stream=0)
    torch.cuda.synchronize()
    _static.taken = free0 - torch.cuda.mem_get_info()[0]      # device memory the call itself took (torch idle in between)
    return d_counts.cpu().numpy(), d_maf.cpu().numpy(), d_rstd.cpu().numpy()


def _check(g, got, label):
    counts, maf, rstd = got
    assert np.array_equal(counts, g["counts"]), label + ": genotype counts"
    assert np.array_equal(maf, g["maf"], equal_nan=True), label + ": maf"
    assert np.array_equal(rstd, g["rstd"], equal_nan=True), label + ": residuals_std"
    assert np.isnan(rstd[3]) and maf[3] == 0.0, label + ": the monomorphic SNP is dropped"


def test_static_scratch_entry_point_matches_on_calls_in_a_row():
    big = _case(1003, 2000, 0.03, seed=71)          # (n_org + 3) / 4 = 251 bytes per SNP: rows start at every alignment
    small = _case(130, 700, 0.0, seed=72)           # fewer rows, other genotypes: the scratch still holds `big`'s
    _check(big, _static(big), "call 1")             # in a fresh process this one loads the kernels and looks the array up
    _check(small, _static(small), "call 2")
    assert _static.taken <= 0, "the entry point must not take device memory"
    _check(big, _static(big), "call 3")
    assert _static.taken <= 0, "the entry point must not take device memory"
    # counts are optional
    _, maf, rstd = _static(small, with_counts=False)
    assert np.array_equal(maf, small["maf"], equal_nan=True) and np.array_equal(rstd, small["rstd"], equal_nan=True)


def test_static_scratch_entry_point_on_a_side_stream():
    g = _case(517, 900, 0.02, seed=73)
    n = g["n_snp"]
    st = torch.cuda.Stream()
    d_packed = torch.from_numpy(g["packed"]).cuda()
    d_counts = torch.empty((n, 3), dtype=torch.int32, device="cuda")
    d_maf = torch.empty((n,), dtype=torch.float64, device="cuda")
    d_rstd = torch.empty((n,), dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    for _ in range(2):                               # back to back on one stream: ordered, no race
        capi.snp_stats_static(d_packed.data_ptr(), n, g["n_org"], MAF, STD, d_counts.data_ptr(), d_maf.data_ptr(),
                              d_rstd.data_ptr(), device=0, stream=st.cuda_stream)
    st.synchronize()
    _check(g, (d_counts.cpu().numpy(), d_maf.cpu().numpy(), d_rstd.cpu().numpy()), "side stream")


@pytest.mark.parametrize("n_org", [1, 3, 4, 5, 9, 16, 127, 129, 1002])
def test_static_scratch_counts_at_every_row_length(n_org):
    """Rows of 1, 2, 3, ... bytes: every mix of row start alignment, neighbour-row bytes in a word and n_org % 4 tail."""
    rng = np.random.default_rng(n_org)
    n = 257
    packed = rng.integers(0, 256, size=(n, (n_org + 3) // 4), dtype=np.uint8)
    _, want = capi.debug_unpack(packed, n_org)
    g = {"packed": packed, "n_org": n_org, "n_snp": n}
    counts, _, _ = _static(g)
    assert np.array_equal(counts, want)


def test_static_scratch_limit_is_refused():
    with pytest.raises(ValueError, match="static scratch limit of 262144"):
        capi.snp_stats_static(256, capi.STATIC_MAX_SNPS + 1, 100, MAF, STD, 0, 256, 256)
    with pytest.raises(ValueError, match="bad arguments"):
        capi.snp_stats_static(0, 10, 100, MAF, STD, 0, 256, 256)
