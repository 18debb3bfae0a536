#!/bin/bash
echo "== in-tree (one patch per CTA)"; python tools/ab_time.py --cases c2,c3,c4,c4b -
echo "== vp persistent, static first patch + pipelined ticket"; python tools/ab_time.py --cases c2,c3,c4,c4b build/vp/libvrt_dyn.so
echo "== vp persistent, static deal"; python tools/ab_time.py --cases c2,c3,c4,c4b build/vp/libvrt_static.so
