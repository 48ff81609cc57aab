"""The reference's OWN integration tests, compiled from the unmodified sources and relinked so that the
gate symbols (bootsNAND/AND/XOR/MUX/...) resolve to libb200tfhe.so while key generation, encryption and
decryption stay in libtfhe (oracle/Makefile target `relink`).  They check every gate output by decrypting
it and print "ERROR!!!" on a mismatch:
  test-addition-boot     (T/test/test-addition-boot.cpp:151-258)  16-bit full_adder_MUX, full_adder,
                         comparison_MUX, 10 trials, n=630 set -- every gate goes through the batch-of-one
                         latency path of the drop-in symbols
  test-gate-bootstrapping (T/test/test-gate-bootstrapping.cpp:56-90)  63-gate NAND trees, 10 trials
"""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RELINK = os.path.join(ROOT, "oracle", "_ref", "relink")

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not os.path.exists(os.path.join(RELINK, "test-addition-boot-gpu")),
                                 reason="oracle/_ref/relink not built (needs /root/reference at build time)")]


@pytest.mark.parametrize("exe,finished", [("test-addition-boot-gpu", 30), ("test-gate-bootstrapping-gpu", 10)])
def test_reference_test_passes_on_gpu_gates(exe, finished):
    r = subprocess.run([os.path.join(RELINK, exe)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:]
    assert "ERROR" not in r.stdout, r.stdout[-2000:]
    assert r.stdout.count("finished bootstrappings") == finished
