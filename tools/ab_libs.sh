# bench several prebuilt variants of the library on one box: bash tools/ab_libs.sh masters-thesis_b200/build/alt_*.so
cp masters-thesis_b200/libb200lstm.so /tmp/base.so
echo base; bash tools/kbench.sh
for f in "$@"; do cp $f masters-thesis_b200/libb200lstm.so; echo $f; bash tools/kbench.sh; done
cp /tmp/base.so masters-thesis_b200/libb200lstm.so
echo base again; bash tools/kbench.sh
