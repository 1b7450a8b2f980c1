"""Generates tests/golden/halton_reference.npz: values of the REFERENCE's own HaltonSampler (samplers/halton.cpp:65-139,
core/lowdiscrepancy.cpp ScrambledRadicalInverse, core/sampler.cpp GlobalSampler::Get1D), taken through
oracle/_ref/ref_harness `halton` from the sampler the reference builds for tests/scenes/box_glass.pbrt:
  q_*      20 000 seeded queries (pixel, sample number, dimension < 1000) -> SampleDimension(GetIndexForSample(n), dim)
  s_*      64 (pixel, sample number) pairs -> the first 96 values of Get1D() after StartPixel / SetSampleNumber
Run in the build container (needs oracle/_ref built from /root/reference):
    python tests/golden/make_golden_halton.py
The queries are regenerated from the seed by the tests (queries()); the file stores them too and the tests check both agree.
"""
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
HARNESS = os.path.join(ROOT, "oracle", "_ref", "ref_harness")
SCENE = os.path.join(ROOT, "tests", "scenes", "box_glass.pbrt")
QUERY = np.dtype([("x", "<i4"), ("y", "<i4"), ("dim", "<i4"), ("count", "<i4"), ("num", "<i8")])
STREAM = 96


def queries():
    rng = np.random.default_rng(2024)
    n = 20000
    q = np.zeros(n, QUERY)
    q["x"], q["y"] = rng.integers(-2, 50, n), rng.integers(-2, 50, n)  # the sample bounds of a 48 x 48 film with a 2-pixel filter
    q["num"] = rng.integers(0, 4096, n)
    q["dim"] = rng.integers(0, 64, n)
    q["dim"][:4000] = rng.integers(0, 1000, 4000)
    q["num"][4000:4400] = rng.integers(0, 2**20, 400)  # Halton indices beyond 2^31 (index = offset + num * 2^7 * 3^4)
    s = np.zeros(64, QUERY)
    s["x"], s["y"] = rng.integers(-2, 50, 64), rng.integers(-2, 50, 64)
    s["num"] = rng.integers(0, 1024, 64)
    s["count"] = STREAM
    return q, s


def main():
    q, s = queries()
    with tempfile.TemporaryDirectory() as tmp:
        qp, op = os.path.join(tmp, "q.bin"), os.path.join(tmp, "o.bin")
        np.concatenate([q, s]).tofile(qp)
        subprocess.check_call([HARNESS, "halton", SCENE, qp, op])
        out = np.fromfile(op, np.float32)
    assert out.size == len(q) + len(s) * STREAM
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "halton_reference.npz"), q=q, s=s,
                        q_value=out[:len(q)], s_value=out[len(q):].reshape(len(s), STREAM))
    print("wrote halton_reference.npz: %d queries, %d streams of %d" % (len(q), len(s), STREAM))


if __name__ == "__main__":
    sys.exit(main())
