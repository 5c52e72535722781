export SWEEP_ONE=1 FSB_GEMV_TMA=1 FSB_GEMV_CPS=1
for kb in 20 40; do for nst in 4 8; do
  [ $((kb*nst)) -gt 170 ] && continue
  echo "stage_kb=$kb nst=$nst: $(FSB_GEMV_STAGE_KB=$kb FSB_GEMV_NST=$nst timeout 100 python tools/gemv_sweep.py head,w13 2>&1 | tail -1)"
done; done
