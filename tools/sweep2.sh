t() { echo "== $1 $2"; DIS_B200_LIB=$1 DIS_WALK_TUNE="$2:$2:$2:$2" python tools/quick_bench.py 3 2>&1 | grep -v "bit-identical" | tail -3; }
for v in n1_16 n1_20 n1_28; do t tools/variants/$v.so 10,24,4; done
t tools/variants/chunk5.so 12,24,4
t tools/variants/chunk8.so 8,24,4
L=optical-flow-using-dense-inverse-search_b200/libdis_b200.so
for s in 10,24,4 10,20,4 10,26,4 16,24,4 10,24,6 10,24,3; do t $L $s; done
