"""On-disk depth formats: byte parity with Benchmark/image_io.py (golden bytes written by the real
reference functions in the build container), the batch normalisation of benchmark.py:760-762, and the async writer."""
import os

import numpy as np
import pytest
import torch

from efficientdepthestimation_b200 import depth_io as D

# image_io.save_raw_float32_image(np.arange(12, dtype=float32).reshape(3, 4) * 0.5)
GOLD_2D = ("0300000004000000050000000400000000000000000000000000003f0000803f0000c03f0000004000002040000040400000604000008040"
           "000090400000a0400000b040")
# image_io.save_raw_float32_image(np.arange(12, dtype=float32).reshape(2, 2, 3) * 0.25 - 1)
GOLD_3D = "0200000002000000150000000c00000000000000000080bf000040bf000000bf000080be000000000000803e0000003f0000403f0000803f0000a03f0000c03f0000e03f"


def test_raw_bytes_match_reference_golden(tmp_path):
    a = np.arange(12, dtype=np.float32).reshape(3, 4) * 0.5
    assert D.raw_float32_bytes(a).hex() == GOLD_2D
    b = np.arange(12, dtype=np.float32).reshape(2, 2, 3) * 0.25 - 1
    assert D.raw_float32_bytes(b).hex() == GOLD_3D
    p = str(tmp_path / "x.raw")
    D.save_raw_float32_image(p, a.astype(np.float64))                 # any dtype is stored as float32
    assert open(p, "rb").read().hex() == GOLD_2D and np.array_equal(D.load_raw_float32_image(p), a)
    D.save_raw_float32_image(p, b)
    assert np.array_equal(D.load_raw_float32_image(p), b)
    with open(p, "r+b") as f:                                          # corrupt pixel_size -> the reference's exception
        f.seek(12)
        f.write((8).to_bytes(8, "little"))
    with pytest.raises(Exception, match="Incompatible pixel_size"):
        D.load_raw_float32_image(p)
    with pytest.raises(ValueError):
        D.raw_float32_bytes(np.zeros(3, np.float32))


def test_normalised_u8_equals_reference_lines():
    g = torch.Generator().manual_seed(0)
    for outputs in (10 * torch.rand(3, 1, 20, 30, generator=g), torch.full((2, 1, 4, 4), 3.0)):      # flat map: no NaN
        n = (outputs - outputs.min()) / (outputs.max() - outputs.min() + torch.finfo(outputs.dtype).tiny)   # benchmark.py:760-761
        ref = np.stack([(255.0 * m).astype("uint8") for m in n.cpu().squeeze(1).numpy()])                   # image_io.py:175-176
        assert np.array_equal(D.normalised_u8(outputs.squeeze(1).numpy()), ref)


@pytest.mark.gpu
def test_async_writer_round_trip(cuda, tmp_path):
    g = torch.Generator().manual_seed(1)
    maps = (10 * torch.rand(5, 1, 57, 76, generator=g)).to(cuda)
    mm = (1000 * maps[:, 0]).to(torch.int32).to(torch.uint16)
    w = D.AsyncDepthWriter(num_threads=2)
    w.write_batch(maps, [str(tmp_path / f"{i:06d}.raw") for i in range(5)], "raw")
    w.write_batch(maps, [str(tmp_path / f"{i:06d}.png") for i in range(5)], "png8")
    w.write_batch(mm, [str(tmp_path / f"mm{i:06d}.png") for i in range(5)], "png16")
    w.cleanup()
    from PIL import Image
    ref8 = D.normalised_u8(maps[:, 0].cpu().numpy())
    for i in range(5):
        assert np.array_equal(D.load_raw_float32_image(str(tmp_path / f"{i:06d}.raw")), maps[i, 0].cpu().numpy())
        assert np.array_equal(np.asarray(Image.open(str(tmp_path / f"{i:06d}.png"))), ref8[i])
        assert np.array_equal(np.asarray(Image.open(str(tmp_path / f"mm{i:06d}.png"))).astype(np.uint16), mm[i].cpu().numpy())
    with pytest.raises(RuntimeError):
        D.AsyncDepthWriter(1).write_batch(torch.zeros(1, 4, 4), ["x"], "raw")
