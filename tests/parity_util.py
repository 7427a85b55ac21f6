"""Measured-distance helpers shared by the GPU parity tests and tools/parity_report.py:
every comparison returns the OBSERVED worst case (so the tests can assert a bar and the
report can print how far inside it the kernels sit)."""
import torch


def u8(t):
    return t.detach().contiguous().view(torch.uint8).cpu()


def ulp_diff(a, b):
    """element-wise |a - b| in fp32 ulps (ordered-integer distance of the bit patterns)."""
    a = a.detach().cpu().contiguous().float().view(torch.int32).long()
    b = b.detach().cpu().contiguous().float().view(torch.int32).long()
    a = torch.where(a < 0, -(a & 0x7FFFFFFF), a)
    b = torch.where(b < 0, -(b & 0x7FFFFFFF), b)
    return (a - b).abs()


def ordered8(code_u8, is_int8=False):
    """8-bit code -> monotone integer (sign-magnitude floats, two's complement int8)."""
    c = code_u8.to(torch.int16)
    if is_int8:
        return torch.where(c >= 128, c - 256, c)
    return torch.where(c >= 128, 127 - c, c)


def code_stats(got, ref, is_int8=False):
    """(number differing, fraction differing, max distance in code steps) for 8-bit codes."""
    got, ref = u8(got), u8(ref)
    assert got.shape == ref.shape, (got.shape, ref.shape)
    neq = got != ref
    n = int(neq.sum())
    if n == 0:
        return 0, 0.0, 0
    d = (ordered8(got[neq], is_int8) - ordered8(ref[neq], is_int8)).abs().max().item()
    return n, n / max(1, got.numel()), int(d)


def nibble_stats(got, ref):
    """packed E2M1 bytes: (number of differing 4-bit codes, fraction, max distance in code steps)."""
    got, ref = u8(got), u8(ref)
    assert got.shape == ref.shape
    lo_g, hi_g, lo_r, hi_r = got & 0xF, got >> 4, ref & 0xF, ref >> 4
    g = torch.stack((lo_g, hi_g), -1).reshape(-1).to(torch.int16)
    r = torch.stack((lo_r, hi_r), -1).reshape(-1).to(torch.int16)
    neq = g != r
    n = int(neq.sum())
    if n == 0:
        return 0, 0.0, 0
    og = torch.where(g >= 8, 7 - g, g)   # sign-magnitude nibble -> monotone
    orr = torch.where(r >= 8, 7 - r, r)
    return n, n / max(1, g.numel()), int((og[neq] - orr[neq]).abs().max())


def scale_stats(got, ref):
    """(number differing, fraction differing, max ulp distance) for fp32 scales."""
    d = ulp_diff(got, ref)
    n = int((d != 0).sum())
    return n, n / max(1, d.numel()), int(d.max()) if d.numel() else 0
