"""Bit-reproducible reductions: ordered split-K weight gradients and ordered column sums, against fp32 PyTorch."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return ((a.float() - b.float()).norm() / (b.float().norm() + 1e-12)).item()


@pytest.mark.parametrize("rows,mw,nw", [(8192, 1024, 1024), (8192, 512, 13), (5000, 256, 400), (300, 1, 256),
                                        (131072, 256, 32)])
def test_wgrad_det_matches_fp32_and_repeats_bit_for_bit(rows, mw, nw):
    from raydp_b200 import ops
    torch.manual_seed(rows + mw)
    g = (torch.randn(rows, (mw + 7) // 8 * 8, device="cuda") * 0.1).to(torch.bfloat16)
    x = torch.randn(rows, (nw + 7) // 8 * 8, device="cuda").to(torch.bfloat16)
    ref = g[:, :mw].float().t() @ x[:, :nw].float()
    ws = torch.empty(8 * mw * nw, device="cuda")
    outs = []
    for _ in range(3):
        out = torch.full((mw, nw), 5.0, device="cuda")          # overwritten, not accumulated
        ops.gemm_wgrad_det(g, x, out, ws, mw=mw, nw=nw)
        outs.append(out)
    assert _rel(outs[0], ref) < 2e-3
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])
    # two-phase form (all layers reduced by one launch) gives the same bits
    out2 = torch.full((mw, nw), 5.0, device="cuda")
    splits = ops.gemm_wgrad_partial(g, x, out2, ws, mw=mw, nw=nw)
    ops.wgrad_reduce_many([ws], [splits], [out2])
    assert torch.equal(out2, outs[0])
    # no scratch: one split, stored directly
    out3 = torch.empty(mw, nw, device="cuda")
    ops.gemm_wgrad_det(g, x, out3, None, mw=mw, nw=nw)
    assert _rel(out3, ref) < 2e-3


@pytest.mark.parametrize("rows", [1, 77, 8192, 40000])
def test_colsum_ordered_matches_fp32_and_repeats_bit_for_bit(rows):
    from raydp_b200 import ops
    torch.manual_seed(rows)
    widths = [1024, 256, 1, 13, 100]
    mats = [torch.randn(rows, (w + 7) // 8 * 8 + 8, device="cuda").to(torch.bfloat16)[:, :(w + 7) // 8 * 8]
            for w in widths]                                                  # pitch != width
    mats.append(torch.randn(rows, 27, 32, device="cuda").to(torch.bfloat16).view(rows, 864)[:, :32])   # slab view
    widths.append(32)
    groups = sum((w + 127) // 128 for w in widths)
    ws = torch.empty(32 * groups * 128, device="cuda")
    counters = torch.zeros(groups, dtype=torch.int32, device="cuda")
    results = []
    for rep in range(3):
        outs = [torch.full((w,), 9.0, device="cuda") for w in widths]
        ops.colsum_ordered(mats, widths, outs, ws if rep < 2 else None, counters if rep < 2 else None)
        results.append(outs)
    for m, w, o in zip(mats, widths, results[0]):
        ref = m[:, :w].float().sum(0)
        assert torch.allclose(o, ref, rtol=1e-4, atol=1e-3 * max(1.0, rows ** 0.5) * 1e-1)
    for a, b in zip(results[0], results[1]):
        assert torch.equal(a, b)
    for a, c in zip(results[0], results[2]):                                  # single-CTA route: same value, other tree
        assert torch.allclose(a, c, rtol=1e-4, atol=1e-2)
    assert (counters == 0).all()
