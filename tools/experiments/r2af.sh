#!/bin/bash
# the UNMODIFIED named-size decks (no sync calls inserted) under STRICT and TRACKED coherence, FAST deposit: wall time
O=$PWD/gpurun_out; mkdir -p $O
export LD_LIBRARY_PATH=/usr/local/cuda/lib64:$LD_LIBRARY_PATH
for deck in harris_c3 lpi_c5; do
  for mode in tracked strict; do
    d=$(mktemp -d); cd $d
    T0=$(date +%s%N)
    env VPIC_B200_COHERENCE=$mode VPIC_B200_DEPOSIT=fast VPIC_B200_REPORT=1 /root/repo/oracle/_ref/deckrun_${deck}_b200 --tpp 1 > $O/r2af_${deck}_$mode.out 2> $O/r2af_${deck}_$mode.err
    rc=$?; T1=$(date +%s%N)
    echo "$deck $mode rc=$rc wall $(( (T1-T0)/1000000 )) ms"; grep "vpic_b200\] device" $O/r2af_${deck}_$mode.err | cut -c1-200
    cd /root/repo; rm -rf $d
  done
done
