"""tcgen05 GEMM vs an fp32 PyTorch reference of the same op."""
import os

import pytest
import torch

from fedrec_b200 import ops
from fedrec_b200.ops import dispatch as D, reference as ref

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _check(M, N, K, epi="bias", block_n=0, out_dtype=torch.bfloat16, m_dev=None, m_scale=1, seed=0, cta_group=0,
           tma_epilogue=0):
    ops.require_native()
    g = torch.Generator(device=DEV).manual_seed(seed)
    x = (torch.randn(M, K, device=DEV, generator=g) * 0.5).bfloat16()
    w = (torch.randn(N, K, device=DEV, generator=g) * (1.0 / K ** 0.5)).bfloat16()
    b = torch.randn(N, device=DEV, generator=g) * 0.1
    r = (torch.randn(M, N, device=DEV, generator=g)).bfloat16() if epi == "resid" else None
    out = torch.full((M, N), 7.0, device=DEV, dtype=out_dtype)
    cnt = torch.tensor([m_dev], device=DEV, dtype=torch.int32) if m_dev is not None else None
    D.linear_bf16(x, w, b, epi, resid=r, out=out, block_n=block_n, m_dev=cnt, m_scale=m_scale, cta_group=cta_group,
                  tma_epilogue=tma_epilogue)
    torch.cuda.synchronize()
    want = ref.linear(x, w, b, {"bias": 0, "gelu": 1, "resid": 2, "tanh": 3}[epi], r)
    rows = M if m_dev is None else min(M, m_dev * m_scale)
    got = out.float()
    err = (got[:rows] - want[:rows]).abs()
    tol = 2e-2 if out_dtype == torch.bfloat16 else 2e-3
    if rows == 0:
        assert (got == 7.0).all()
        return
    if float(err.max()) > tol * max(1.0, float(want.abs().max())):
        os.makedirs("gpurun_out", exist_ok=True)
        bad = (err > tol).nonzero()
        with open("gpurun_out/gemm_diag.txt", "a") as f:
            f.write(f"M={M} N={N} K={K} epi={epi} bn={block_n} max_err={float(err.max())} n_bad={bad.shape[0]}\n")
            f.write(f" first bad idx: {bad[:16].tolist()}\n got {got[:4, :8].tolist()}\n want {want[:4, :8].tolist()}\n")
            f.write(f" bad rows hist(mod 128): {torch.bincount(bad[:, 0] % 128, minlength=128).tolist()}\n")
            f.write(f" bad cols hist(mod 64): {torch.bincount(bad[:, 1] % 64, minlength=64).tolist()}\n")
    torch.testing.assert_close(got[:rows], want[:rows], atol=tol, rtol=tol)
    if rows < M:
        assert (got[rows + 127:] == 7.0).all()                     # rows past the dynamic M are untouched


@pytest.mark.parametrize("block_n", [64, 128, 256])
def test_square_tiles(block_n):
    _check(256, 256, 128, block_n=block_n)


def test_single_k_block():
    _check(128, 64, 64, block_n=64)


@pytest.mark.parametrize("M,N,K,epi", [
    (300, 768, 768, "bias"), (1000, 2304, 768, "bias"), (513, 3072, 768, "gelu"), (257, 768, 3072, "resid"),
    (2500, 1200, 400, "bias"), (2500, 200, 400, "tanh"), (82500, 768, 768, "bias"), (4099, 384, 768, "tanh"),
])
def test_model_shapes(M, N, K, epi):
    _check(M, N, K, epi)


@pytest.mark.parametrize("cta_group", [1, 2])
@pytest.mark.parametrize("tma_epilogue", [-1, 1])
@pytest.mark.parametrize("M,N,K,epi", [(4500, 768, 768, "resid"), (8200, 3072, 768, "gelu"), (5000, 768, 3072, "resid"),
                                       (300, 2304, 768, "bias"), (256, 256, 64, "bias"), (6000, 400, 192, "tanh")])
def test_kernel_variants(cta_group, tma_epilogue, M, N, K, epi):
    """Every (CTA pair | single CTA) x (staged TMA | direct) epilogue variant on the wide tile."""
    _check(M, N, K, epi, block_n=256, cta_group=cta_group, tma_epilogue=tma_epilogue)


@pytest.mark.parametrize("tma_epilogue", [-1, 1])
@pytest.mark.parametrize("M,N,K,epi", [(4500, 768, 768, "resid"), (5000, 768, 3072, "resid"), (4200, 3072, 768, "gelu"),
                                       (300, 768, 768, "bias"), (4100, 400, 192, "tanh")])
def test_narrow_pair_tile(tma_epilogue, M, N, K, epi):
    """192-column tiles on CTA pairs (three of the four epilogue column groups active; N = 400 clips the last tile)."""
    _check(M, N, K, epi, block_n=192, cta_group=2, tma_epilogue=tma_epilogue)
    _check(6000, 768, 768, "resid", block_n=192, m_dev=133, m_scale=30, cta_group=2, tma_epilogue=tma_epilogue)


@pytest.mark.parametrize("cta_group", [1, 2])
def test_variants_with_device_side_m(cta_group):
    _check(6000, 768, 768, "resid", block_n=256, m_dev=133, m_scale=30, cta_group=cta_group, tma_epilogue=1)
    _check(6000, 768, 768, "gelu", block_n=256, m_dev=7, m_scale=30, cta_group=cta_group, tma_epilogue=1)


def test_fp32_output_small_m():
    _check(130, 400, 768, out_dtype=torch.float32)
    _check(2750, 400, 768, out_dtype=torch.float32)


def test_dynamic_m_from_device():
    _check(3000, 768, 768, m_dev=37, m_scale=30)
    _check(3000, 768, 768, m_dev=0, m_scale=30)
    _check(3000, 768, 768, m_dev=1000, m_scale=30)                 # larger than M: clamps


def test_many_waves_persistent():
    _check(148 * 128 * 3 + 5, 256, 192, block_n=256)


@pytest.mark.parametrize("M,N,K,epi,tma", [
    (26880, 768, 768, "resid", 1),      # 315 tiles on 74 pairs: the 19 tiles of the last round are cut in two
    (26880, 768, 3072, "resid", -1),    # same, direct epilogue (the FFN2 variant)
    (26880, 3072, 768, "gelu", 1),      # 1260 tiles: 2 left over
    (2000, 768, 768, "bias", 1),        # 24 tiles <= half the pairs: every tile is split
    (2100, 1000, 768, "bias", 1),       # ragged N: the right half of the last column tile is partly (or wholly) past N
    (26880, 2304, 768, "bias", 1),      # 945 tiles: last round of 57 > half the pairs, nothing is split
])
def test_tail_split_of_a_short_last_round(M, N, K, epi, tma):
    """CTA-pair kernels deal tiles round-robin; a last round that fills at most half of the pairs is run as 256 x 128
    half tiles (N = 128 MMAs, 64-row B boxes). Same results as the fp32 reference, with and without a device-side M."""
    _check(M, N, K, epi, block_n=256, cta_group=2, tma_epilogue=tma, seed=3)
    _check(M, N, K, epi, block_n=256, cta_group=2, tma_epilogue=tma, seed=4, m_dev=M // 30 - 3, m_scale=30)
