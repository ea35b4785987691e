mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 --print-limit 20 python __graft_entry__.py smoke > gpurun_out/memcheck_smoke.log 2>&1; echo "memcheck smoke rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt; tail -15 gpurun_out/memcheck_smoke.log
