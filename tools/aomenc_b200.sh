#!/bin/sh
# BASELINE configs[0] (aomenc --cpu-used=6 --good, synthetic CIF clip, 10 frames, 1 thread) with the stock
# AVX2 dispatch and with the `_b200` RTCD flavour bound at run time (oracle/_ref/libb200install.so):
# wall time of both runs, the number of symbols bound and a byte comparison of the two .ivf files.
# usage: tools/aomenc_b200.sh [groups mask, default 15 = fwd txfm | quant | block error | inv txfm]
G=${1:-15}
D=$(mktemp -d)
python baseline/gen_y4m.py $D/cif.y4m
A="oracle/_ref/aomenc_avx2 --cpu-used=6 --good --threads=1 --limit=10 -q"
s=$(date +%s.%N); $A -o $D/stock.ivf $D/cif.y4m; e=$(date +%s.%N)
echo "stock avx2: $(python -c "print(round($e - $s, 2))") s, $(stat -c %s $D/stock.ivf) bytes"
s=$(date +%s.%N); LD_PRELOAD=oracle/_ref/libb200install.so B200_RTCD_GROUPS=$G $A -o $D/b200.ivf $D/cif.y4m; e=$(date +%s.%N)
echo "b200 flavour (groups $G): $(python -c "print(round($e - $s, 2))") s, $(stat -c %s $D/b200.ivf) bytes"
cmp $D/stock.ivf $D/b200.ivf && echo "byte-identical"
rm -rf $D
