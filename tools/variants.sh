for v in "FSB_X=1" "FSB_GEMV_NST=6" "FSB_GEMV_NST=10" "FSB_PDL=0" "FSB_GEMV_CPS=2 FSB_GEMV_NST=4"; do
  env $v timeout 100 python bench.py --no-cpu --steps 1 --frames 128 --e2e-steps 0 > gpurun_out/v.json 2> gpurun_out/v.err; echo "$v rc=$? $(python -c "import json;d=json.load(open('gpurun_out/v.json'));print(d['mean_ms_per_frame'], d['roofline']['frac'])" 2>/dev/null) $(grep -m1 -E 'Error|error' gpurun_out/v.err | cut -c1-120)"
done
