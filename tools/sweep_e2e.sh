cd /root/repo
for s in "4" "1" "3" "6" "8" "4,3,2,1" "5,4,3,2,1" "6,5,4,3,2,1" "8,4,2,1" "3,3,2,1" "4,4,3,2,1" "16,12,9,6,4,2"; do
  echo "== DIS_OUT_SLICES=$s"
  DIS_OUT_SLICES=$s python tools/e2e_bench.py 10 2>&1 | tail -1
done
