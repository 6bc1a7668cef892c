for S in 1 2 4; do for LS in 1; do
  BB_STBL_SPLIT=$S BB_STBL_LSPLIT=$LS timeout 200 python bench.py --workload c5s --steps 5 --warmup 3 --no-cpu --no-ess --no-sub --e2e-steps 1 > gpurun_out/sw.log 2>&1
  echo "split $S lambda-split $LS: $(python scripts/show_bench.py gpurun_out/sw.log | grep HEADLINE | cut -c1-70)"
done; done
