"""One erode() step bracketed by cudaProfilerStart/Stop, for `ncu --profile-from-start off` (needs a GPU).
usage: ncu ... python tools/ncu_step.py LG WARMUP_STEPS [KIND]"""
import ctypes, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from veloren_b200 import synth
from veloren_b200.erosion import ErosionContext
lg, warm = int(sys.argv[1]), int(sys.argv[2])
kind = int(sys.argv[3]) if len(sys.argv) > 3 else 0
rt = ctypes.CDLL("libcudart.so")
c = ErosionContext(lg, lg); c.configure(synth.generate(lg, lg, kind=kind))
c.run_device(warm)
rt.cudaProfilerStart()
c.step()
rt.cudaProfilerStop()
print("profiled step", warm, "n_gs", int(c.scalar("n_gs")), "lakes", int(c.scalar("n_lakes")))
