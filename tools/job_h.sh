#!/bin/bash
for l in - build/mb10/libvrt.so build/mb11/libvrt.so build/mb13/libvrt.so build/mb14/libvrt.so build/mb16/libvrt.so; do echo "== $l"; python tools/ab_time.py --cases c2,c3,c4 $l; done
