#!/bin/bash
for l in build/vq/libvrt.so; do echo "== $l"; VRT_LIBRARY=$l timeout 250 python tools/queue_probe.py; done
