"""PCG32 streams as Dr.Jit lays them out (reference: ext/drjit/include/drjit/random.h:44-73, 131-133), on torch
tensors. `mi.PCG32(size=n)` is n independent generators, generator j seeded with initstate + j and initseq + j;
the reference's sampling scripts draw their inputs from it (sunsky-testing/sampling_test.py:30-32,
src/emitters/tests/test_sunsky.py:215-218), so BASELINE's "fixed RNG inputs" are these floats. Input generation
only: nothing here is on the measured path."""
import torch

PCG32_DEFAULT_STATE = 0x853C49E6748FEA9B
PCG32_DEFAULT_STREAM = 0xDA3E39CB94B95BDB
_MULT = 0x5851F42D4C957F2D


def _i64(v):
    """Python int (mod 2^64) -> the same bits as a signed 64-bit value."""
    v &= (1 << 64) - 1
    return v - (1 << 64) if v >= (1 << 63) else v


def _lsr(x, k):
    """Logical shift right of int64 tensors."""
    return (x >> k) & ((1 << (64 - k)) - 1)


class PCG32:
    def __init__(self, size, initstate=PCG32_DEFAULT_STATE, initseq=PCG32_DEFAULT_STREAM, device="cpu"):
        j = torch.arange(size, dtype=torch.int64, device=device)
        # seed(initstate + j, initseq + j) (random.h:57-65)
        self.state = torch.zeros(size, dtype=torch.int64, device=device)
        self.inc = ((j + _i64(initseq)) << 1) | 1
        self.next_uint32()
        self.state = self.state + (j + _i64(initstate))
        self.next_uint32()

    def next_uint32(self):
        old = self.state
        self.state = old * _i64(_MULT) + self.inc                      # wraps mod 2^64
        xorshifted = _lsr(_lsr(old, 18) ^ old, 27) & 0xFFFFFFFF
        rot = _lsr(old, 59)
        return ((xorshifted >> rot) | (xorshifted << ((-rot) & 31))) & 0xFFFFFFFF

    def next_float32(self):
        """(bits >> 9) | 0x3f800000 reinterpreted, minus 1 (random.h:131-133): in [0, 1)."""
        bits = ((self.next_uint32() >> 9) | 0x3F800000).to(torch.int32)
        return bits.view(torch.float32) - 1.0


def floats(n, draws=1, stream_offset=0, device="cpu"):
    """(draws, n) float32: draw k of generator stream_offset + j."""
    rng = PCG32(n, PCG32_DEFAULT_STATE + stream_offset, PCG32_DEFAULT_STREAM + stream_offset, device=device)
    return torch.stack([rng.next_float32() for _ in range(draws)])


def point2_pairs(n, stream_offset=0, device="cpu"):
    """`PCG32(size=2n).next_float32()` reshaped (n, 2) as sampling_test.py:30-32 does -> planar (2, n):
    sample i = (float of generator 2i, float of generator 2i + 1)."""
    f = floats(2 * n, 1, stream_offset, device)[0]
    return f.view(n, 2).t().contiguous()


def configs1_inputs(n, device, rank=0):
    """SURVEY.md 8(d) cfg 2: generator j of PCG32(size = 2n) gives one float; (u1, u2) = (float[2i], float[2i + 1]);
    cos(theta) = u1, phi = 2 pi u2, wo = (sin cos, sin sin, cos), si.wi = -wo; wavelengths = 360 + 360 *
    sample_shifted<4>(u3) with u3 from the next block of n generators (core/math.h:408-431). Ranks use disjoint
    stream blocks. Returns (wi (3, n), wo (3, n), wavelengths (4, n)) as torch tensors on `device`."""
    import numpy as np
    base = rank * 3 * n
    u = point2_pairs(n, stream_offset=base, device=device)
    u3 = floats(n, 1, stream_offset=base + 2 * n, device=device)[0]
    ct = u[0]; st = torch.sqrt(torch.clamp(1 - ct * ct, min=0)); ph = (2 * np.pi) * u[1]
    wo = torch.stack([st * torch.cos(ph), st * torch.sin(ph), ct]).contiguous()
    shifted = u3[None, :] + torch.arange(4, device=device, dtype=torch.float32)[:, None] / 4
    shifted = torch.where(shifted >= 1, shifted - 1, shifted)
    # 360 + 360 * (1 - 2^-24) rounds to 720.0f, where the reference gathers one bin past its tables
    wav = (360.0 + 360.0 * shifted).clamp_(max=719.99).contiguous()
    return (-wo).contiguous(), wo, wav
