for tool in memcheck racecheck synccheck initcheck; do
  timeout 900 compute-sanitizer --tool $tool --error-exitcode 7 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r01c_sanitizer_$tool.log 2>&1; echo "$tool rc=$?"
  tail -4 gpurun_out/r01c_sanitizer_$tool.log
done
