#!/usr/bin/env python3
"""Mint the golden fixture of BASELINE configs[0] (independent bin streams through the range coder) from the
UNMODIFIED reference coder: oracle/_ref/ref_harness `arith` drives arithmetic_code.h exactly as
test/arithmetic_code.cpp:93-110 does.  Output: tests/golden/streams.json
  cases[]: {digit_bits, model, nstreams, nbins, bits (hex of np.packbits of the stream-major 0/1 array), coded[] (hex per stream)}
Streams: Bernoulli p(1) in {0.5, 0.1, 0.01, 0.001}, plus carry-heavy patterns (long runs of ones
push `low` to the top of the interval), all-zero, all-one, alternating; sizes chosen ragged (not multiples of 8).
"""
import json, os, subprocess, tempfile
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.path.join(ROOT, "oracle", "_ref", "ref_harness")


def make_bits(nstreams, nbins, seed):
    rng = np.random.RandomState(seed)
    rows = []
    probs = [0.5, 0.1, 0.01, 0.001, 0.9, 0.999]
    for s in range(nstreams):
        kind = s % 10
        if kind < 6:
            rows.append((rng.random_sample(nbins) < probs[kind]).astype(np.uint8))
        elif kind == 6:
            rows.append(np.ones(nbins, np.uint8))
        elif kind == 7:
            rows.append(np.zeros(nbins, np.uint8))
        elif kind == 8:
            rows.append((np.arange(nbins) % 2).astype(np.uint8))
        else:   # long runs of ones broken by single zeros at random places
            r = np.ones(nbins, np.uint8); r[rng.randint(0, max(nbins, 1), size=max(nbins // 97, 1))] = 0 if nbins else 0
            rows.append(r)
    return np.stack(rows) if nbins else np.zeros((nstreams, 0), np.uint8)


cases = []
for digit_bits in (16, 8):
    for model in (0, 1):
        for nstreams, nbins, seed in ((20, 1003, 1), (3, 0, 2), (5, 1, 3), (4, 7, 4), (4, 8, 5), (4, 9, 6), (2, 20011, 7)):
            bits = make_bits(nstreams, nbins, seed * 100 + digit_bits + model)
            with tempfile.NamedTemporaryFile() as f:
                f.write(bits.tobytes()); f.flush()
                out = subprocess.run([REF, "arith", str(digit_bits), str(model), str(nstreams), str(nbins), f.name],
                                     check=True, stdout=subprocess.PIPE).stdout.decode()
            coded = [l[2:].strip() for l in out.splitlines() if l.startswith("A")]
            assert len(coded) == nstreams
            cases.append(dict(digit_bits=digit_bits, model=model, nstreams=nstreams, nbins=nbins,
                              bits=np.packbits(bits.reshape(-1)).tobytes().hex(), coded=coded))
json.dump(dict(source="oracle/_ref/ref_harness arith (reference arithmetic_code.h, test/arithmetic_code.cpp:93-110)", cases=cases),
          open(os.path.join(HERE, "streams.json"), "w"))
print(len(cases), "cases,", os.path.getsize(os.path.join(HERE, "streams.json")), "bytes")
