"""Writes tests/golden/test_spirv.json: the 32-bit words of the reference's checked-in compiled shader
/root/reference/RayTracing/Assets/Shaders/Compiled/Ray/Test.spirv (1,524 bytes, sha256 recorded in SURVEY.md §8c),
so that the pin test can run where /root/reference does not exist (the GPU box).  Run in the build container:

    python tests/golden/make_spirv_golden.py

A compiled-module word list is a reference-held golden vector, not reference source; tests/test_spirv_pin.py re-reads
the original file whenever it is present and checks that this fixture is still identical to it.
"""
import json
import os
import struct
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import spirv_eval as E  # noqa: E402

SRC = "/root/reference/RayTracing/Assets/Shaders/Compiled/Ray/Test.spirv"


def main():
    blob = open(SRC, "rb").read()
    digest = E.sha256(blob)
    assert digest == E.TEST_SPIRV_SHA256, digest
    words = struct.unpack(f"<{len(blob) // 4}I", blob)
    out = {"source": "RayTracing/Assets/Shaders/Compiled/Ray/Test.spirv", "bytes": len(blob), "sha256": digest,
           "words_hex": [f"{w:08x}" for w in words]}
    with open(os.path.join(HERE, "test_spirv.json"), "w") as f:
        json.dump(out, f, indent=0)
    print(f"wrote {len(words)} words, sha256 {digest}")


if __name__ == "__main__":
    main()
