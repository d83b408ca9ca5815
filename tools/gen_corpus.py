#!/usr/bin/env python3
"""gen_corpus.py <dir> <first> <count>: writes the bench corpus clips (bench.py clip_params) to <dir>."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
d, first, count = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
os.makedirs(d, exist_ok=True)
bench.gen_corpus(d, first, count, keep_files=True)
