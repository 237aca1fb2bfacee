#!/bin/bash
# where the start-up of the resident deck programs goes (first-touch uploads, page-locking)
O=$PWD/gpurun_out; mkdir -p $O
export LD_LIBRARY_PATH=/usr/local/cuda/lib64:$LD_LIBRARY_PATH
for deck in lpi_c5 harris_c3; do
  d=$(mktemp -d); cd $d
  T0=$(date +%s.%N); env VPIC_B200_COHERENCE=resident VPIC_B200_DEPOSIT=fast VPIC_B200_REPORT=1 VPIC_B200_TRACE_CALLS=5 VPIC_B200_TRACE_TRANSFERS=1 \
     /root/repo/oracle/_ref/deckrun_${deck}_res_b200 --tpp 1 > $O/r2ac_$deck.out 2> $O/r2ac_$deck.err
  echo "$deck rc=$? wall $(echo "$(date +%s.%N) - $T0" | bc) s"; grep -c "h2d" $O/r2ac_$deck.err; grep "vpic_b200\] [a-z_]*:" $O/r2ac_$deck.err | sort -t: -k2 -n -r | head -12
  cd /root/repo
done
