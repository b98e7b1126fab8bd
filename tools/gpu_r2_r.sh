for v in "" shade8 spawn3 both; do
  if [ -z "$v" ]; then L=""; else L="lib=build/variants/$v/librt_b200.so"; fi
  timeout 200 python tools/frame_time.py c4 $L reps=4 -- world=8 reps=4 2>&1 | grep -v "^$"
done
