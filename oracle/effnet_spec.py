"""EfficientNet-B0..B7 architecture tables of efficientnet-pytorch 0.6.3.

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).

The reference imports the third-party package ``efficientnet-pytorch==0.6.3``
(``/root/reference/requirements.txt:8``; call sites ``ReSIDE/models/modules.py:8,166-198``
and ``ReSIDE/models/lasinger2019.py:4,57-91``).  The package source is not under
/root/reference and cannot be installed offline, so its published algorithm is
restated here from SURVEY.md Appendix A ("parity unpinned" against upstream).
"""
import math
from dataclasses import dataclass
from typing import List, Tuple

# name -> (width multiplier, depth multiplier, nominal resolution, dropout)
MODEL_PARAMS = {
    "efficientnet-b0": (1.0, 1.0, 224, 0.2),
    "efficientnet-b1": (1.0, 1.1, 240, 0.2),
    "efficientnet-b2": (1.1, 1.2, 260, 0.3),
    "efficientnet-b3": (1.2, 1.4, 300, 0.3),
    "efficientnet-b4": (1.4, 1.8, 380, 0.4),
    "efficientnet-b5": (1.6, 2.2, 456, 0.4),
    "efficientnet-b6": (1.8, 2.6, 528, 0.5),
    "efficientnet-b7": (2.0, 3.1, 600, 0.5),
}

# (repeats, kernel, stride, expand ratio, in filters, out filters, se ratio) of the base net.
BASE_STAGES = [
    (1, 3, 1, 1, 32, 16, 0.25),
    (2, 3, 2, 6, 16, 24, 0.25),
    (2, 5, 2, 6, 24, 40, 0.25),
    (3, 3, 2, 6, 40, 80, 0.25),
    (3, 5, 1, 6, 80, 112, 0.25),
    (4, 5, 2, 6, 112, 192, 0.25),
    (1, 3, 1, 6, 192, 320, 0.25),
]

BN_EPS = 1e-3          # batch_norm_epsilon
BN_MOMENTUM = 0.01     # 1 - batch_norm_momentum(0.99)
DEPTH_DIVISOR = 8

# block index splits used by the reference (modules.py:168-181, lasinger2019.py:59-72)
SPLITS = {
    "efficientnet-b0": [0, 3, 5, 8],
    "efficientnet-b1": [0, 5, 8, 16],
    "efficientnet-b3": [0, 5, 8, 18],
    "efficientnet-b4": [0, 6, 10, 22],
    "efficientnet-b5": [0, 8, 13, 27],
    "efficientnet-b6": [0, 9, 15, 31],
    "efficientnet-b7": [0, 11, 18, 38],
}


@dataclass(frozen=True)
class BlockSpec:
    kernel: int
    stride: int
    expand: int
    cin: int
    cout: int
    stage_first: bool      # stride was decoded as the list [s] -> the id-skip test fails

    @property
    def cexp(self) -> int:
        return self.cin * self.expand

    @property
    def csq(self) -> int:
        return max(1, int(self.cin * 0.25))

    @property
    def skip(self) -> bool:
        # `id_skip and stride == 1 and in == out`; `[1] == 1` is False for stage-first blocks
        return (not self.stage_first) and self.stride == 1 and self.cin == self.cout


def round_filters(filters: int, width: float) -> int:
    filters *= width
    new = max(DEPTH_DIVISOR, int(filters + DEPTH_DIVISOR / 2) // DEPTH_DIVISOR * DEPTH_DIVISOR)
    if new < 0.9 * filters:
        new += DEPTH_DIVISOR
    return int(new)


def round_repeats(repeats: int, depth: float) -> int:
    return int(math.ceil(depth * repeats))


def block_specs(name: str) -> Tuple[int, List[BlockSpec], int]:
    """-> (stem width, per-block specs, nominal image size)."""
    width, depth, res, _ = MODEL_PARAMS[name]
    blocks: List[BlockSpec] = []
    for (r, k, s, e, i, o, _se) in BASE_STAGES:
        i, o, r = round_filters(i, width), round_filters(o, width), round_repeats(r, depth)
        blocks.append(BlockSpec(k, s, e, i, o, True))
        for _ in range(r - 1):
            blocks.append(BlockSpec(k, 1, e, o, o, False))
    return round_filters(32, width), blocks, res


def static_same_pad(kernel: int, stride: int, image_size: int) -> Tuple[int, int]:
    """0.6.x static 'same' padding: computed once from the NOMINAL image size for every
    layer (Appendix A.4).  -> (before, after) for both axes (top==left, bottom==right)."""
    out = math.ceil(image_size / stride)
    pad = max((out - 1) * stride + (kernel - 1) + 1 - image_size, 0)
    return pad // 2, pad - pad // 2


def split_indices(name: str) -> List[int]:
    if name not in SPLITS:
        raise RuntimeError(f"Unsupported encoder network '{name}'.")
    _, blocks, _ = block_specs(name)
    return SPLITS[name] + [len(blocks)]
