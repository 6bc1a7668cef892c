for W in c5s c4w; do for S in 4 6 8; do for LS in 1; do
  BB_STBL_SPLIT=$S BB_STBL_LSPLIT=$LS timeout 200 python bench.py --workload $W --steps 5 --warmup 3 --no-cpu --no-ess --no-sub --e2e-steps 1 > gpurun_out/sw.log 2>&1
  echo "$W split $S lambda-split $LS: $(python scripts/show_bench.py gpurun_out/sw.log | grep HEADLINE | cut -c1-70)"
done; done; done
