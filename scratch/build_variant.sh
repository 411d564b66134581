#!/bin/bash
# build_variant.sh NAME [make args, e.g. EXTRA=-DJPLG_MIX_FMA PROF=1] -> scratch/libjplg_NAME.so (experiment builds for same-box A/B runs)
set -e
name=$1; shift
root=$(cd "$(dirname "$0")/.." && pwd)
tmp=/tmp/jplg_variant_$name
rm -rf $tmp && mkdir -p $tmp/jplspatial_b200 $tmp/include
cp -r $root/jplspatial_b200/csrc $tmp/jplspatial_b200/csrc
cp $root/include/*.h $tmp/include/
rm -f $tmp/jplspatial_b200/csrc/*.o
make -C $tmp/jplspatial_b200/csrc -j8 OUT=$root/scratch/libjplg_$name.so "$@" > $tmp/make.log 2>&1 || { tail -20 $tmp/make.log; exit 1; }
ls -la $root/scratch/libjplg_$name.so
