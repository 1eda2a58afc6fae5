#!/bin/bash
# SASS opcode counts per object of librtseg.so: the instructions that prove which hardware paths the kernels use
# (UTCHMMA = tcgen05.mma, UTMALDG / UTMASTG = TMA tensor load / store, LDTM = tcgen05.ld, UTCBAR = tcgen05.commit,
#  SYNCS = mbarrier, F2FP.RELU = fused ReLU in the bf16 conversion).   bash tools/sass_summary.sh > profiles/sass_opcodes_rNN.txt
cd "$(dirname "$0")/../realtime-semantic-segmentation-pytorch_b200/build" || exit 1
echo "object            UTCHMMA UTMALDG UTMASTG  LDTM UTCBAR SYNCS F2FP.RELU  FFMA2"
for f in conv_tc pair_tc dwpw_tc stem_tc dw_tma train train_dw fuse_ops optim conv_dw conv_dense stem head ohem runtime; do
  [ -f $f.o ] || continue
  cuobjdump -sass $f.o 2>/dev/null | awk -v n=$f.o '/UTCHMMA/{a++} /UTMALDG/{b++} /UTMASTG/{c++} /LDTM/{d++} /UTCBAR/{e++} /SYNCS/{f++} /F2FP.*RELU/{g++} /FFMA2/{h++} END{printf "%-16s %8d %7d %7d %5d %6d %5d %9d %6d\n",n,a,b,c,d,e,f,g,h}'
done
