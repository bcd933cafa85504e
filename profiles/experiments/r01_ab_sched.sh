export AB_ROUNDS=1
B=bwa-mem2-arm_b200/ab
echo "mode0"; python tools/ab_bench.py $B/base.so
echo "mode1"; BSW_CUDA_BIN_MODE=1 python tools/ab_bench.py $B/base.so
echo "mode2"; BSW_CUDA_BIN_MODE=2 python tools/ab_bench.py $B/base.so
echo "persist nocap"; python tools/ab_bench.py $B/persist.so
for caps in "9:1,8:2,7:3" "9:1,8:3,7:4" "9:1,8:2,7:3,6:4" "9:1,8:4,7:5,6:6" "9:1,8:1,7:2,6:2,5:3"; do
  echo "persist caps $caps"; BSW_CUDA_CAPS=$caps python tools/ab_bench.py $B/persist.so
done
