t() { echo "== $1 $2"; DIS_B200_LIB=$1 DIS_WALK_TUNE="$2:$2:$2:$2" python tools/quick_bench.py 3 2>&1 | grep -v "bit-identical" | tail -3; }
L=optical-flow-using-dense-inverse-search_b200/libdis_b200.so
for s in 8,24,4 8,24,2; do t $L $s; done
for s in 10,24,2 10,24,4; do t tools/variants/chunk6.so $s; done
for s in 16,24,2 16,20,2; do t tools/variants/chunk4.so $s; done
for s in 20,24,2 20,16,2; do t tools/variants/chunk3.so $s; done
for s in 32,24,2 32,16,2 32,8,0; do t tools/variants/chunk2.so $s; done
