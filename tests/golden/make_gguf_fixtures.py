#!/usr/bin/env python
"""Generates the GGUF fixtures of tests/golden (run in the authoring container, where oracle/_ref exists):

  reference_testfile.gguf  byte-for-byte what the reference's own TestGGUFGenerator.generateTestFile() emits
                           (src/nativeTest/kotlin/ai/solace/llamakotlin/gguf/TestGGUFGenerator.kt:13-82): v3, 3 KV pairs,
                           two F32 tensors — transcribed field by field, the Kotlin code cannot run here.
  ggml_written.gguf        written by the reference tree's own C writer (gguf_write_to_file, ggml/src/ggml.c, through
                           oracle/_ref): KV pairs of 5 value kinds incl. a string array, and four small tensors
                           (F32, Q4_0, Q4_1, Q8_0) quantized by ggml's quantize_row_q*_ref => ggml type ids, ggml nibble
                           order, ggml's own alignment padding.
  ggml_written.json        what the file holds (names, shapes, sha256 of every payload, the F32 source of the quantized
                           tensors' seeds) for the reader tests.
"""
import hashlib
import json
import os
import struct
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))


def kotlin_test_file() -> bytes:
    def s(x):
        b = x.encode()
        return struct.pack("<Q", len(b)) + b
    buf = b"GGUF" + struct.pack("<IQQ", 3, 2, 3)
    buf += s("general.architecture") + struct.pack("<I", 8) + s("test")
    buf += s("general.name") + struct.pack("<I", 8) + s("test-model")
    buf += s("general.alignment") + struct.pack("<I", 10) + struct.pack("<Q", 32)
    buf += s("weight.0") + struct.pack("<IQQIQ", 2, 2, 2, 0, 0)
    buf += s("weight.1") + struct.pack("<IQQIQ", 2, 3, 3, 0, 16)
    buf += b"\0" * ((32 - len(buf) % 32) % 32)
    buf += struct.pack("<4f", 1, 2, 3, 4) + struct.pack("<9f", 1, 0, 0, 0, 1, 0, 0, 0, 1)
    return buf


def main():
    from oracle import ref as R
    assert R.build() is not None, "needs /root/reference (oracle/_ref)"
    with open(os.path.join(HERE, "reference_testfile.gguf"), "wb") as f:
        f.write(kotlin_test_file())
    K, M = 128, 24
    tensors, meta = [], {}
    for i, (name, t) in enumerate((("token_embd.weight", 0), ("blk.0.attn_q.weight", 2), ("blk.0.ffn_up.weight", 3),
                                   ("blk.0.ffn_down.weight", 6))):
        w = np.random.default_rng(7000 + i).normal(0, 0.02, K * M).astype(np.float32)
        payload = w if t == 0 else R.quantize(t, w)
        tensors.append((name, t, K, M, payload))
        meta[name] = {"type": t, "ne": [K, M], "seed": 7000 + i, "nbytes": int(payload.nbytes),
                      "sha256": hashlib.sha256(payload.tobytes()).hexdigest()}
    R.gguf_write(os.path.join(HERE, "ggml_written.gguf"), tensors, alignment=32)
    with open(os.path.join(HERE, "ggml_written.json"), "w") as f:
        json.dump({"alignment": 32, "tensors": meta}, f, indent=1)
    print("wrote", os.listdir(HERE))


if __name__ == "__main__":
    main()
