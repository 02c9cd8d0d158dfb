#!/bin/bash
# Builds video_style_transfer_b200/libvst_b200_alt.so with extra -D flags on the named sources (A/B timing
# of two builds in one GPU call: VST_LIB_PATH=.../libvst_b200_alt.so).  usage: build_alt.sh "<flags>" file.cu ...
set -e
FLAGS="$1"; shift
CS=video_style_transfer_b200/csrc
mkdir -p /tmp/vst_alt
OBJS=""
for f in $CS/*.cu; do
  b=$(basename $f .cu)
  if [[ " $* " == *" $b.cu "* ]]; then
    nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC --expt-relaxed-constexpr $FLAGS -c $f -o /tmp/vst_alt/$b.o
    OBJS="$OBJS /tmp/vst_alt/$b.o"
  else
    OBJS="$OBJS $CS/$b.o"
  fi
done
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o video_style_transfer_b200/libvst_b200_alt.so $OBJS
ls -la video_style_transfer_b200/libvst_b200_alt.so
