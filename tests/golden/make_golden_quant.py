"""Generates tests/golden/quantize_sbhalf.npz by running the REFERENCE's own rowwise quantizer
(third_party/fbgemm/src/QuantUtils.cc, compiled unmodified into oracle/_ref/libfbgemm_quant_ref.so by
oracle/Makefile): FloatOrHalfToFusedNBitRowwiseQuantizedSBHalfRef (:499) and the dispatcher (:574) that the
torch op FloatToFusedNBitRowwiseQuantizedSBHalf calls (fbgemm_gpu/src/quantize_ops_cpu.cpp).

Run here (the container with /root/reference):   python tests/golden/make_golden_quant.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import c_oracle  # noqa: E402


def cases():
    rng = np.random.default_rng(17)
    out = {}
    for name, (rows, cols, scale) in {"d128": (96, 128, 1.0), "d64_small": (40, 64, 1e-3), "d20": (33, 20, 7.0),
                                      "d256_large": (17, 256, 3e3), "d3": (9, 3, 1.0)}.items():
        w = (rng.standard_normal((rows, cols)) * scale).astype(np.float32)
        w[0] = 0.25            # constant row: range 0 -> scale 1
        w[1] = 0.0
        w[2, 0] = 65504.0 * 4  # minimum/scale overflow paths of the fp16 casts
        w[3] = np.linspace(-1, 1, cols, dtype=np.float32)  # exact ties after scaling
        out[name] = w
    return out


def main():
    assert c_oracle.quant_ref_lib() is not None, "build oracle/_ref first (make -C oracle)"
    blob = {}
    for name, w in cases().items():
        ref = c_oracle.ref_quantize_sbhalf(w, dispatch=False)
        disp = c_oracle.ref_quantize_sbhalf(w, dispatch=True)
        assert np.array_equal(ref, disp), f"{name}: the reference's AVX2 kernel and its scalar definition differ"
        blob[name + ".w"], blob[name + ".q"] = w, ref
    np.savez_compressed(os.path.join(HERE, "quantize_sbhalf.npz"), **blob)
    print("wrote", len(blob) // 2, "cases")


if __name__ == "__main__":
    main()
