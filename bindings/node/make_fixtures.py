"""Fixtures for bindings/node/b200tfhe.test.js, written by the UNMODIFIED reference (oracle/_ref, built by
oracle/Makefile).  Run in the build container:   python bindings/node/make_fixtures.py
Small key (n = 8, otherwise the 80-bit shape) so that the files stay small: cloud.key (1.5 MB), a/b/c.ct,
bits.json, lwe_key.json, and for eoc's string entry points eoc.json = base64 strings of two Msize-encoded
integers (eoc-tfhe-run.cpp:282-300), of the reference's lweAddTo / lweSubTo results, and of two 8-bit
integers encrypted bit by bit together with the cloud key in base64."""
import base64
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import reflib  # noqa: E402
from test_eoc_entry_points import EocRef  # noqa: E402  (libtfhe calls eoc-tfhe-run.cpp makes)

out = os.path.join(HERE, "fixtures")
os.makedirs(out, exist_ok=True)
ref = reflib.RefLib()
rk = ref.keygen_custom(n=8, l=2, Bgbit=10, seed=2025)
rk.export_cloud(os.path.join(out, "cloud.key"))
rng = np.random.default_rng(5)
bits = {}
for tag in "abc":
    b = rng.integers(0, 2, size=12).astype(np.int32)
    bits[tag] = b.tolist()
    open(os.path.join(out, tag + ".ct"), "wb").write(rk.export_samples(rk.encrypt_bits(b)))
json.dump(bits, open(os.path.join(out, "bits.json"), "w"))
json.dump(rk.lwe_key().tolist(), open(os.path.join(out, "lwe_key.json"), "w"))

e = EocRef(rk)
s15, s27 = e.encrypt(15), e.encrypt(27)
x, y, nb = 0xB7, 0x5C, 8
xb = rk.encrypt_bits([(x >> i) & 1 for i in range(nb)])
yb = rk.encrypt_bits([(y >> i) & 1 for i in range(nb)])
json.dump({
    "ct15": e.to_b64(s15), "ct27": e.to_b64(s27),
    "sum": e.to_b64(e.combine(s15, s27, sub=False)), "diff": e.to_b64(e.combine(s27, s15, sub=True)),
    "cloud_key": base64.b64encode(open(os.path.join(out, "cloud.key"), "rb").read()).decode(),
    "x": x, "y": y, "nb_bits": nb, "n": rk.n,
    "x_bits": base64.b64encode(rk.export_samples(xb)).decode(), "y_bits": base64.b64encode(rk.export_samples(yb)).decode(),
}, open(os.path.join(out, "eoc.json"), "w"))
print("wrote", sorted(os.listdir(out)))
