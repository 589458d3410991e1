#!/bin/bash
# per-layer timing of the 3x3 stride-1 layers of ResNet-50 / 224 with different halo boxes (probing script)
run() { python scripts/profile_conv.py $@ 2>&1 | tail -1; }
for kind in 0 1; do
echo "== 56x56 64->64 kind $kind"; echo -n "off      "; RN_TC_HALO=0 run 128 56 56 64 64 3 1 1 $kind 5
for box in "8 16" "14 9" "28 4" "56 2" "4 28" "7 18"; do set -- $box; echo -n "box $1x$2  "; RN_TC_HALO=1 RN_TC_HALO_BW=$1 RN_TC_HALO_BH=$2 run 128 56 56 64 64 3 1 1 $kind 5; done
echo "== 28x28 128->128 kind $kind"; echo -n "off      "; RN_TC_HALO=0 run 128 28 28 128 128 3 1 1 $kind 5
for pair in 0 1; do for box in "28 4" "4 28" "14 9" "7 18"; do set -- $box; echo -n "pair $pair box $1x$2  "; RN_TC_PAIR=$pair RN_TC_HALO=1 RN_TC_HALO_BW=$1 RN_TC_HALO_BH=$2 run 128 28 28 128 128 3 1 1 $kind 5; done; done
echo "== 14x14 256->256 kind $kind"; echo -n "off      "; RN_TC_HALO=0 run 128 14 14 256 256 3 1 1 $kind 5
for box in "14 9" "14 7" "7 14"; do set -- $box; echo -n "box $1x$2  "; RN_TC_HALO=1 RN_TC_HALO_BW=$1 RN_TC_HALO_BH=$2 run 128 14 14 256 256 3 1 1 $kind 5; done
done
