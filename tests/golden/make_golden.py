#!/usr/bin/env python3
"""Mint the golden fixtures from the UNMODIFIED reference (run in the build container, needs
oracle/_ref/ built by oracle/build_ref.sh from /root/reference):
  <name>.h264      synthetic clip (avrecode-ms_b200/bin/avrgen, parameters below)
  <name>.avrec     output of the reference's `recode compress` on it
  <name>.trace.gz  hook-event trace of the reference's libavcodec-hooks parse (oracle/ref_harness.cpp)
Committed so that tests can pin the oracle and the product without the reference present.
"""
import gzip, os, subprocess, sys
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
GEN = os.path.join(ROOT, "avrecode-ms_b200", "bin", "avrgen")
REF = os.path.join(ROOT, "oracle", "_ref")
CLIPS = {
    "ip_basic": "seed=101 mb_w=6 mb_h=4 frames=3",
    "multislice_multiref": "seed=102 mb_w=8 mb_h=5 frames=4 slices_per_frame=3 num_ref=3",
    "dct8x8": "seed=103 mb_w=6 mb_h=4 frames=3 transform_8x8=1",
    "dense_full_blocks": "seed=104 mb_w=5 mb_h=3 frames=2 p_sig=250 p_last=2 p_cbf=250 p_cbp=250 transform_8x8=1",
    "skippy_idr": "seed=105 mb_w=10 mb_h=6 frames=5 p_skip=200 idr_period=2 qp=40 cabac_init_idc=2 slices_per_frame=2",
    # round 2: B slices (every picture a reference, so frame_num keeps changing: inside the reference's own domain)
    "b_slices": "seed=106 mb_w=8 mb_h=5 frames=5 p_bslice=190 num_ref=3 slices_per_frame=2 cabac_init_idc=1",
    "b_direct_multiref": "seed=107 mb_w=7 mb_h=6 frames=6 p_bslice=220 p_direct=140 p_skip=90 num_ref=4",
}
for name, params in CLIPS.items():
    clip = os.path.join(HERE, name + ".h264")
    subprocess.run([GEN, clip] + params.split(), check=True, stderr=subprocess.DEVNULL)
    subprocess.run([os.path.join(REF, "recode"), "compress", clip, os.path.join(HERE, name + ".avrec")], check=True,
                   stderr=subprocess.DEVNULL)
    tr = subprocess.run([os.path.join(REF, "ref_harness"), "trace", clip], check=True, stdout=subprocess.PIPE,
                        stderr=subprocess.DEVNULL).stdout
    with gzip.GzipFile(os.path.join(HERE, name + ".trace.gz"), "wb", mtime=0) as f:
        f.write(tr)
    print(name, os.path.getsize(clip), os.path.getsize(os.path.join(HERE, name + ".avrec")), len(tr))
