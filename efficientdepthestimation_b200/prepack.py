"""Weight prepacking for the sm_100a kernels (host logic, torch tensor ops only).

Reference layout: conv weights OIHW fp32, BatchNorm as (weight, bias, running_mean,
running_var) -- SURVEY.md 8b.  Kernel layout (include/ede.h):
  * ``w_direct``  fp32 [kh*kw][cin_p][cout_p]           (SIMT implicit GEMM, exact fp32 path)
  * ``w_umma``    bf16 [kh*kw][cout_p16][cin_p64] K-major (tcgen05 path, TMA-loaded B operand)
  * ``scale/shift`` fp32 [cout_p]: BatchNorm folded as per-channel affine applied to the fp32
    accumulator (NOT folded into bf16 weights: one rounding less per layer).
Channel padding: every activation view has a multiple-of-8 lane count; concatenations are
segments at 8-aligned offsets.  ``in_map[j]`` gives, for input lane j of the view, the source
channel of the reference conv (or -1 for a zero lane); ``out_map`` likewise for output lanes.
"""
from typing import List, Optional, Sequence, Tuple

import torch


def ceil8(c: int) -> int:
    return (c + 7) // 8 * 8


def ceil_to(c: int, m: int) -> int:
    return (c + m - 1) // m * m


def identity_map(c: int, c_p: Optional[int] = None) -> List[int]:
    c_p = ceil8(c) if c_p is None else c_p
    return list(range(c)) + [-1] * (c_p - c)


def scatter_weight(weight: torch.Tensor, in_map: Sequence[int], out_map: Sequence[int]) -> torch.Tensor:
    """OIHW reference weight -> dense [cout_p][cin_p][kh][kw] fp32 in kernel lane order."""
    w = weight.detach().float()
    dev = w.device
    kh, kw = w.shape[2], w.shape[3]
    full = torch.zeros(len(out_map), len(in_map), kh, kw, dtype=torch.float32, device=dev)
    oi = torch.tensor([i for i, s in enumerate(out_map) if s >= 0], dtype=torch.long, device=dev)
    os_ = torch.tensor([s for s in out_map if s >= 0], dtype=torch.long, device=dev)
    ii = torch.tensor([i for i, s in enumerate(in_map) if s >= 0], dtype=torch.long, device=dev)
    is_ = torch.tensor([s for s in in_map if s >= 0], dtype=torch.long, device=dev)
    sub = w.index_select(0, os_).index_select(1, is_)
    full[oi.unsqueeze(1), ii.unsqueeze(0)] = sub
    return full


def pack_direct(full: torch.Tensor) -> torch.Tensor:
    """[cout_p][cin_p][kh][kw] -> fp32 [kh*kw][cin_p][cout_p]."""
    co, ci, kh, kw = full.shape
    return full.permute(2, 3, 1, 0).reshape(kh * kw, ci, co).contiguous()


def pack_umma(full: torch.Tensor, passes: int = 1) -> Tuple[torch.Tensor, int, int]:
    """[cout_p][cin_p][kh][kw] -> bf16 [passes*kh*kw][cout_p16][cin_p64] (K-major rows).
    passes=2 appends the low halves: taps [0,kh*kw) hold bf16(w), taps [kh*kw, 2*kh*kw) hold
    bf16(w - bf16(w)), so that accumulating both gives 16 mantissa bits of the weight."""
    co, ci, kh, kw = full.shape
    co_p, ci_p = ceil_to(co, 16), ceil_to(ci, 64)
    taps = full.permute(2, 3, 0, 1).reshape(kh * kw, co, ci)
    out = torch.zeros(passes * kh * kw, co_p, ci_p, dtype=torch.bfloat16, device=full.device)
    hi = taps.to(torch.bfloat16)
    out[:kh * kw, :co, :ci] = hi
    if passes == 2:
        out[kh * kw:, :co, :ci] = (taps - hi.float()).to(torch.bfloat16)
    return out.contiguous(), co_p, ci_p


def pack_umma_taps(full: torch.Tensor, passes: int, g: int) -> Tuple[torch.Tensor, int, int]:
    """Tap-packed weights for the operand-swapped tcgen05 form with `g` (2 or 4) taps of a kernel row per MMA
    (include/ede.h, EdeConv.w_pack_g): [cout_p][cin_p][kh][kw] -> bf16 [passes*kh*ceil(kw/g)][128][cin_p64] with
    row j*(128/g)+c of tile (ky, gx) = w[c][.][ky][gx*g + j]; rows of taps beyond kw and of channels beyond
    cout_p stay zero.  passes=2 appends the bf16 low halves as in pack_umma."""
    co, ci, kh, kw = full.shape
    cg = 128 // g
    if co > cg:
        raise ValueError(f"pack_umma_taps: {co} output lanes do not fit {g} taps in 128 rows")
    kws, ci_p = (kw + g - 1) // g, ceil_to(ci, 64)
    tiles = torch.zeros(kh * kws, 128, ci, dtype=torch.float32, device=full.device)
    for ky in range(kh):
        for kx in range(kw):
            gx, j = divmod(kx, g)
            tiles[ky * kws + gx, j * cg:j * cg + co] = full[:, :, ky, kx]
    out = torch.zeros(passes * kh * kws, 128, ci_p, dtype=torch.bfloat16, device=full.device)
    hi = tiles.to(torch.bfloat16)
    out[:kh * kws, :, :ci] = hi
    if passes == 2:
        out[kh * kws:, :, :ci] = (tiles - hi.float()).to(torch.bfloat16)
    return out.contiguous(), 128, ci_p


PACK_N = 16      # include/ede.h EDE_PACK_N: kernel-row taps on the UMMA N axis


def pack_umma_ntaps(full: torch.Tensor, passes: int, g: int) -> Tuple[torch.Tensor, int, int]:
    """Tap-packed weights for the pixels-on-M tcgen05 form with `g` (2..5) taps of a kernel row on the N axis
    (EdeConv.w_pack_g = EDE_PACK_N | g): [cout_p][cin_p][kh][kw] -> bf16 [passes*kh*ceil(kw/g)][g*cout_p16][cin_p64]
    with row j*cout_p16+c of tile (ky, gx) = w[c][.][ky][gx*g + j]; rows of taps beyond kw and of channels beyond
    cout_p stay zero.  passes=2 appends the bf16 low halves as in pack_umma."""
    co, ci, kh, kw = full.shape
    slot = ceil_to(co, 16)
    if g * slot > 256:
        raise ValueError(f"pack_umma_ntaps: {g} taps of {slot} output lanes exceed 256 UMMA columns")
    kws, ci_p = (kw + g - 1) // g, ceil_to(ci, 64)
    tiles = torch.zeros(kh * kws, g * slot, ci, dtype=torch.float32, device=full.device)
    for ky in range(kh):
        for kx in range(kw):
            gx, j = divmod(kx, g)
            tiles[ky * kws + gx, j * slot:j * slot + co] = full[:, :, ky, kx]
    out = torch.zeros(passes * kh * kws, g * slot, ci_p, dtype=torch.bfloat16, device=full.device)
    hi = tiles.to(torch.bfloat16)
    out[:kh * kws, :, :ci] = hi
    if passes == 2:
        out[kh * kws:, :, :ci] = (tiles - hi.float()).to(torch.bfloat16)
    return out.contiguous(), g * slot, ci_p


def fold_bn(cout: int, bn: Optional[dict], conv_bias: Optional[torch.Tensor], device) -> Tuple[torch.Tensor, torch.Tensor]:
    """-> (scale, shift) fp32 [cout] with  bn(conv(x)+b) == scale*conv(x) + shift.
    bn: dict(weight, bias, mean, var, eps) or None."""
    if bn is None:
        scale = torch.ones(cout, dtype=torch.float32, device=device)
        shift = torch.zeros(cout, dtype=torch.float32, device=device)
    else:
        inv = torch.rsqrt(bn["var"].detach().double() + bn["eps"])
        scale64 = bn["weight"].detach().double() * inv
        shift64 = bn["bias"].detach().double() - bn["mean"].detach().double() * scale64
        scale, shift = scale64.float().to(device), shift64.float().to(device)
    if conv_bias is not None:
        shift = shift + scale * conv_bias.detach().float().to(device)
    return scale, shift


def place_lanes(vec: torch.Tensor, out_map: Sequence[int], fill: float = 0.0) -> torch.Tensor:
    """Per-channel vector of the reference -> kernel lane order (pad lanes = fill)."""
    dev = vec.device
    out = torch.full((len(out_map),), fill, dtype=torch.float32, device=dev)
    oi = torch.tensor([i for i, s in enumerate(out_map) if s >= 0], dtype=torch.long, device=dev)
    os_ = torch.tensor([s for s in out_map if s >= 0], dtype=torch.long, device=dev)
    out[oi] = vec.detach().float().index_select(0, os_)
    return out
