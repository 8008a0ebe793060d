import sys, time
sys.path.insert(0, "/root/repo")
import flex_gpu_scheduler_b200 as fgs
s = fgs.config_synth(2)
eng = fgs.Engine(device=0)
eng.set_plugins(fgs.config_plugins(2))
eng.upload_nodes(fgs.Packer(2).nodes(s.snapshot))
for pin in (True, False):
    t = fgs.Packer(2, pin=pin).pods(s.pending)
    for _ in range(3):
        eng.evaluate(t)
    ts = []
    for _ in range(10):
        t0 = time.perf_counter(); eng.evaluate(t); ts.append(time.perf_counter() - t0)
    tm = eng.timings()
    print("pinned" if pin else "pageable", "wall %.3f ms" % (sorted(ts)[5] * 1e3), {k: round(tm[k], 3) for k in ("h2d_ms", "reduce_ms", "emit_ms", "d2h_ms")})
