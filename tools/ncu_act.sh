OUT=gpurun_out/${1:-r01f}; mkdir -p $OUT
for cfg in "384 13792 f32" "24 220672 bf16"; do
  tag=$(echo $cfg | tr ' ' '_')
  C1="python tools/run_one.py act $cfg"
  timeout 120 $C1 > $OUT/plain_$tag.log 2>&1 &&
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:act1d -s 2 -c 1 -f -o $OUT/prof_act_$tag $C1 > $OUT/ncu_$tag.log 2>&1
  echo "ncu $tag rc=$?"
done
