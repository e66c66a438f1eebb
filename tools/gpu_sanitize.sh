# compute-sanitizer over small solves that reach every kernel variant
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
{
echo "== plain"; timeout 300 python tools/sanitize_cases.py
for tool in memcheck racecheck synccheck; do
  echo "== $tool"
  timeout 1500 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_cases.py dense sparse batch 2>&1 | grep -v "^dense\|^sparse\|^batch" | tail -25
done
echo "== memcheck pairs"; timeout 1200 compute-sanitizer --tool memcheck --print-limit 20 python tools/sanitize_cases.py pairs 2>&1 | tail -8
echo "== racecheck pairs"; timeout 1200 compute-sanitizer --tool racecheck --print-limit 20 python tools/sanitize_cases.py pairs 2>&1 | tail -12
} > gpurun_out/sanitize.log 2>&1
cat gpurun_out/sanitize.log
