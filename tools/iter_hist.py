import sys, os, time
sys.path[:0] = [os.path.dirname(os.path.dirname(os.path.abspath(__file__))), os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests")]
import _golden, lut_ldpc_b200 as L
cfg = sys.argv[1] if len(sys.argv) > 1 else "c5_n64800"
d = L.Decoder(_golden.bundle_path(cfg)); d.set_exit_conditions(d.max_iters, 1, 1)
for snr in [float(x) for x in (sys.argv[2] if len(sys.argv) > 2 else "1.1,1.3,1.5,2.0").split(",")]:
    t0 = time.time(); c = d.simulate_snr(snr, int(sys.argv[3]) if len(sys.argv) > 3 else 4096, seed=1, batch_frames=4096); dt = time.time() - t0
    h = list(c.iter_hist); st = d.last_call_stats()
    print("snr", snr, c.as_dict(), "wall %.3f dev_ms %.1f" % (dt, st["device_ms"]))
    print("  hist", {i: int(v) for i, v in enumerate(h) if v})
