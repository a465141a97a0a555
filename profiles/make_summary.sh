# builds profiles/r01/ncu_summary_final.md from the reports profiles/ncu_r01_final.sh left in gpurun_out/
out=profiles/r01/ncu_summary_final.md
cp gpurun_out/launches_final.csv profiles/r01/launches_final.csv
{
echo '# ncu summaries, round 1 final kernels (12 MP 4000x3000, -t 1048576 -q 0.25; `bash profiles/ncu_r01_final.sh`, then `bash profiles/make_summary.sh`)'
echo
echo 'Raw reports stay in gpurun_out/ (scratch); this file and launches_final.csv are the tracked summaries.'
echo 'Kernels as of this capture: transform with dp2a colour conversion / byte row pass / column-at-a-time DCT (80 registers);'
echo 'entropy tiles of 128 blocks, zero-padded survivor lists, unconditional window ORs.  ncu flushes the caches before every'
echo 'kernel: k_entropy_prepare reads its 36 MB from DRAM here, from L2 in a real fit (the transform has just written them).'
echo
echo '## launch list (ncu --metrics gpu__time_duration.sum --clock-control none; bench.py --steps 1 --warmup 1 --batch 8)'
echo '```'
python profiles/launch_summary.py gpurun_out/launches_final.csv
echo '```'
for k in pass transform prepare emit resize; do
  echo; echo "## $k (ncu --set full --clock-control none --import-source on; first captured launch)"; echo '```'
  ncu -i gpurun_out/prof_${k}_final.ncu-rep --page raw --csv 2>/dev/null | python profiles/ncu_keys.py 2>/dev/null | awk '/^-- /{n++} n<2'
  echo '```'
  if [ $k != resize ]; then
    echo; echo "### $k: dynamic SASS mix"; echo '```'
    ncu -i gpurun_out/prof_${k}_final.ncu-rep --page source --csv 2>/dev/null | python profiles/sass_mix.py 12
    echo '```'
  fi
done
} > $out 2>&1
grep -n "gpu__time_duration\|dram__bytes\|^-- \|inst_executed.sum\|issue_active\|fmaheavy" $out
