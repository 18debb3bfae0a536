#!/bin/bash
echo "== in-tree, sun field auto"; python tools/ab_time.py --cases c2,c3,c4 - 
echo "== in-tree, sun field off"; VRT_SUN_FIELD=0 python tools/ab_time.py --cases c2,c3,c4 -
echo "== old face rule, sun field auto"; python tools/ab_time.py --cases c2,c3,c4 build/vofr/libvrt.so
echo "== old face rule, sun field off"; VRT_SUN_FIELD=0 python tools/ab_time.py --cases c2,c3,c4 build/vofr/libvrt.so
echo "== base 6b8e830"; python tools/ab_time.py --cases c2,c3,c4 build/base/libvrt.so
