OUT=gpurun_out/${1:-r01c}; mkdir -p $OUT
for cfg in "24 220672 11 1 1" "96 55168 11 1 1" "24 220672 3 1 0"; do
  tag=$(echo $cfg | tr ' ' '_')
  C1="python tools/run_one.py conv $cfg"
  timeout 120 $C1 > $OUT/plain_$tag.log 2>&1 &&
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_umma -s 2 -c 1 -f -o $OUT/prof_conv_$tag $C1 > $OUT/ncu_$tag.log 2>&1
  echo "ncu $tag rc=$?"
done
