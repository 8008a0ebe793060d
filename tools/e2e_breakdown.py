"""Where the time of one fgs_evaluate call goes (C2): python tools/e2e_breakdown.py"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import flex_gpu_scheduler_b200 as fgs
s = fgs.config_synth(2)
eng = fgs.Engine(device=0)
eng.set_plugins(fgs.config_plugins(2))
eng.upload_nodes(fgs.Packer(2).nodes(s.snapshot))
t = fgs.Packer(2).pods(s.pending)
for k in range(6):
    t0 = time.perf_counter()
    r = eng.evaluate(t)
    wall = (time.perf_counter() - t0) * 1e3
    tm = eng.timings()
    print("wall %.3f ms  stage+h2d %.3f  reduce %.3f  emit %.3f  finalize+d2h %.3f" % (wall, tm["h2d_ms"], tm["reduce_ms"], tm["emit_ms"], tm["d2h_ms"]))
