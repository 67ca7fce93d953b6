"""Pins the oracle's sampler restatement against the compiled reference (oracle/_ref): in
"compiled order" mode (the operation order gcc -O3 -ffast-math produced, see mlt_oracle.c) the
nearest, bilinear and bicubic samplers must reproduce interp*_b32/_b64 of libmltplus.so with ZERO
mismatches; the literal-source mode stays within +-1 LSB.  CPU only."""
import numpy as np
import pytest

import oracle
import refmlt

pytestmark = pytest.mark.skipif(not refmlt.available(), reason="oracle/_ref not built")


def samples(bits, n, seed):
    rng = np.random.default_rng(seed)
    w = h = 9
    dt = np.uint8 if bits == 8 else np.uint16
    hi = 256 if bits == 8 else 65536
    for i in range(n):
        s = rng.integers(0, hi, w * h * 4).astype(dt)
        d = rng.integers(0, hi, 4).astype(dt)
        if i % 2:
            s[3::4] = hi - 1
        if i % 3 == 0:
            d[3] = hi - 1
        if i % 11 == 0:
            s[3::4] = 0
            d[3] = 0  # 0/0 alpha -> NaN -> 0
        yield s, d, w, h, float(rng.uniform(0, w - 1)), float(rng.uniform(0, h - 1)), float(
            rng.choice([1.0, 0.8, 0.5, 0.33, 0.0])), int(i % 5 == 0)


@pytest.mark.parametrize("bits", [8, 16])
@pytest.mark.parametrize("kind", [0, 1, 2])
def test_compiled_order_is_bit_exact_with_reference(bits, kind):
    oracle.set_compiled_order(1)
    try:
        bad = 0
        for s, d, w, h, x, y, o, atop in samples(bits, 4000, 7 + bits + kind):
            if kind == 2:
                x, y = min(max(x, 0.01), w - 1.0), min(max(y, 0.01), h - 1.0)
            r = refmlt.interp(bits, kind, s, w, h, x, y, o, d, atop)
            g = oracle.interp_px(bits, kind, s, w, h, x, y, o, d, atop)
            bad += int((r != g).sum())
        assert bad == 0
    finally:
        oracle.set_compiled_order(1)


@pytest.mark.parametrize("bits", [8, 16])
@pytest.mark.parametrize("kind", [0, 1, 2])
@pytest.mark.parametrize("mode", [0, 1])
def test_all_modes_within_1lsb(bits, kind, mode):
    oracle.set_compiled_order(mode)
    try:
        worst = 0
        for s, d, w, h, x, y, o, atop in samples(bits, 1500, 99 + bits + kind):
            if kind == 2:
                x, y = min(max(x, 0.01), w - 1.0), min(max(y, 0.01), h - 1.0)
            r = refmlt.interp(bits, kind, s, w, h, x, y, o, d, atop).astype(int)
            g = oracle.interp_px(bits, kind, s, w, h, x, y, o, d, atop).astype(int)
            worst = max(worst, int(np.abs(r - g).max()))
        assert worst <= 1
    finally:
        oracle.set_compiled_order(1)
