#!/bin/bash
# The library's own memcheck (compute-sanitizer is closed on this GPU pool): run smoke() and the small-scene parity tests
# against the RS_CHECKED build, in which every indexed access of the wave state and every stack push is bounds-checked on
# the device and a render with a failed check returns RS_ERR_CUDA.  Then the negative control: the first 32 paths of
# every wave also store their record one buffer too far — the same render MUST be refused with the count of caught stores.
#   make VARIANT=_chk TUNE=-DRS_CHECKED=1 && bash tests/tools/checked_run.sh > gpurun_out/rXX_checked.txt
export RAYSNAIL_B200_LIB=$PWD/raysnail_b200/libraysnail_b200_chk.so
echo "== smoke() on the checked build"
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
echo "== parity tests on the checked build"
python -m pytest tests -m gpu -q -x -k "${CHECKED_K:-primary_hit or single_bounce or pixel_mask or wave_size or passes_driver or list_and_trees or feature_scene or mis_integrator or rs_multi}" 2>&1 | tail -3
echo "== negative control: injected out-of-range stores, the render must be refused"
RAYSNAIL_B200_CHECK_SELFTEST=1 python - <<'PY' 2>&1 | tail -3
from raysnail_b200 import Renderer, Scene, make_params
s = Scene.builtin("rtow_final", 7)
r = Renderer(0)
r.upload(s)
try:
    r.render(s.camera(96, 60), make_params(samples=16))
    print("NEGATIVE CONTROL FAILED: no violation reported")
except Exception as e:
    print("negative control ok:", e)
PY
