"""Parity gate P4: our kernels against the REFERENCE's own CUDA kernels, compiled
unmodified for sm_100a (oracle/Makefile -> oracle/_ref/libref_aqlm.so), on the same B200 and the
same tensors.  Skipped when the library was not built (it needs /root/reference at build time)."""
import pytest
import torch

import vllm_pq_b200 as P
from oracle import pq_oracle as O
from oracle import ref_cuda as R
from vllm_pq_b200 import _lib

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not R.available(), reason="oracle/_ref/libref_aqlm.so not built")]

FORMATS = [(16, 8, 1), (8, 8, 2)]          # the two formats the reference has kernels for


def dev(c):
    return {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in c.items()}


@pytest.mark.parametrize("fmt", FORMATS)
@pytest.mark.parametrize("parts,k", [([256], 512), ([4096], 4096), ([512, 512], 1024)])
def test_dequant_matches_reference_cuda(fmt, parts, k):
    """Single partition: bit-exact.  2x8 with several partitions: the reference CUDA dequant reads
    codebooks [p, p+1] instead of [2p, 2p+1] (gemm_kernels.cu:582-585 vs :462-464);
    partition 0 still agrees bit for bit, and ours agrees with the reference's torch path."""
    b, d, m = fmt
    c = dev(O.make_case(parts, k, b, d, m, t=1, dtype=torch.float16, seed=21))
    ours = P.ops.pq_dequant(c["codes"], c["codebooks"], parts)
    ref = R.aqlm_dequant(c["codes"], c["codebooks"], parts)
    torch.cuda.synchronize()
    if len(parts) == 1 or m == 1:
        assert torch.equal(ours.view(torch.int16), ref.view(torch.int16))
    else:
        # ... except in the CTA that straddles the partition boundary: the reference stages ONE
        # shared-memory codebook per CTA from per-warp pointers (gemm_kernels.cu:246-275), so the
        # last <= 16 rows (thread_m, :350-358) of a partition whose size is not a multiple of the
        # CTA's row count see a mix of both partitions' codebooks.  Compare below that.
        n0 = parts[0] - 16
        assert torch.equal(ours[:n0].view(torch.int16), ref[:n0].view(torch.int16))
        oracle = O.dequantize_partitions(c["codes"].cpu(), c["codebooks"].cpu(), parts, None)
        assert torch.equal(ours.cpu().view(torch.int16), oracle.view(torch.int16))


@pytest.mark.parametrize("fmt", FORMATS)
@pytest.mark.parametrize("t", [1, 3, 6])
@pytest.mark.parametrize("parts,k", [([4096], 4096), ([1024, 512, 512], 2048)])
@pytest.mark.parametrize("bias", [False, True])
def test_gemv_matches_reference_cuda(fmt, t, parts, k, bias):
    """aqlm_gemm (Code1x16MatVec / Code2x8MatVec, one launch per token) vs pq::gemm on the same
    tensors.  The reference accumulates each 8-group in fp16 (gemm_kernels.cu:84-87,166-170), we
    accumulate in fp32, so the gate is a tolerance: both must be within the P3 bound of the fp64
    contraction, and ours must not be further from it than the reference kernel is."""
    b, d, m = fmt
    c = O.make_case(parts, k, b, d, m, t=t, dtype=torch.float16, seed=5 + t, bias=bias)
    g = dev(c)
    ours = P.ops.pq_gemm(g["x"], g["codes"], g["codebooks"], g["scales"], parts, g["bias"])
    ref = R.aqlm_gemm(g["x"], g["codes"], g["codebooks"], g["scales"], parts, g["bias"])
    torch.cuda.synchronize()
    exact = O.linear_fp32(c["x"], c["codes"], c["codebooks"], c["scales"], parts, c["bias"])
    e_ours = O.error_report(ours.cpu(), exact)
    assert e_ours["rms_rel"] <= 2e-3, e_ours
    # Code2x8MatVec stages ONE shared-memory codebook per CTA from per-warp pointers
    # (gemm_kernels.cu:114-141) and its rows per CTA follow the SM count (:331-338, <= 16): on a
    # 148-SM part CTAs straddle the partition boundaries and up to 16 rows either side of one see
    # the wrong partition's codebooks.  Those rows are left out of the comparison WITH THE
    # REFERENCE KERNEL only (ours is checked on every row above).
    keep = torch.ones(sum(parts), dtype=torch.bool)
    if m > 1 and len(parts) > 1:
        edge = 0
        for p_rows in parts[:-1]:
            edge += p_rows
            keep[max(0, edge - 16):edge + 16] = False
        # ... and the last CTA: its warps beyond the last row never advance their codebook pointer
        # (gemm_kernels.cu:117-126) but still take part in the staging loop
        keep[-16:] = False
    e_ours_k = O.error_report(ours.cpu()[:, keep], exact[:, keep])
    e_ref = O.error_report(ref.cpu()[:, keep], exact[:, keep])
    assert e_ref["rms_rel"] <= 4e-3, e_ref
    assert e_ours_k["rms_rel"] <= e_ref["rms_rel"] * 1.5 + 1e-6, (e_ours_k, e_ref)
    torch.testing.assert_close(ours.float()[:, keep.cuda()], ref.float()[:, keep.cuda()], rtol=2e-2,
                               atol=2e-2 * float(exact.abs().max()))


@pytest.mark.parametrize("fmt", FORMATS)
def test_prefill_matches_reference_dequant_gemm(fmt):
    """optimized_dequantize_gemm over the reference CUDA dequant (aqlm.py:132-156) vs our tcgen05 GEMM."""
    b, d, m = fmt
    parts, k, t = [2048], 2048, 256
    c = O.make_case(parts, k, b, d, m, t=t, dtype=torch.float16, seed=8)
    g = dev(c)
    ours = P.ops.pq_gemm(g["x"], g["codes"], g["codebooks"], g["scales"], parts, None, algo=_lib.ALGO_GEMM_TC)
    ref = R.optimized_dequantize_gemm(g["x"], g["codes"], g["codebooks"], g["scales"], parts, None)
    torch.cuda.synchronize()
    exact = O.linear_fp32(c["x"], c["codes"], c["codebooks"], c["scales"], parts, None)
    e_ours, e_ref = O.error_report(ours.cpu(), exact), O.error_report(ref.cpu(), exact)
    assert e_ours["rms_rel"] <= 2e-3 and e_ours["rms_rel"] <= e_ref["rms_rel"] * 1.5 + 1e-6, (e_ours, e_ref)
