#!/usr/bin/env python
"""One-screen summary of bench.py JSON lines.  Usage: python tools/bench_extract.py file.json [...]"""
import json
import sys

for f in sys.argv[1:]:
    for l in open(f):
        l = l.strip()
        if not l.startswith('{'):
            continue
        d = json.loads(l)
        rf = d.get('roofline') or {}
        c5, tm = rf.get('c5') or {}, rf.get('tamols') or {}
        print(f"{f}: n_gpus {d.get('n_gpus')} impl {d.get('impl', 'cuda')}")
        print(f"  C2 device p50 {d.get('p50_latency_ms')} ms, value {d.get('value'):.4g}, e2e p50 {d['e2e'].get('p50_latency_ms')} ms, "
              f"class {d['e2e'].get('class_call_p50_latency_ms')}, frac {rf.get('frac')}")
        cc = d['e2e'].get('c_client')
        if cc:
            print(f"  C2 e2e from a plain C client: p50 {cc['p50_latency_ms'] * 1e3:.2f} us, p99 {cc['p99_latency_ms'] * 1e3:.2f} us")
        dt = d.get('device_timing')
        if dt:
            print(f"  device timing floor {dt['event_pair_floor_p50_ms'] * 1e3:.2f} us, C2 solve minus floor {dt['solve_p50_minus_floor_ms'] * 1e3:.2f} us")
        if c5:
            print(f"  C5 launch {c5.get('launch_ms'):.4f} ms step {c5.get('ms_per_step'):.4f} ms frac {c5.get('frac'):.4f} "
                  f"(whole step {c5.get('frac_whole_step'):.4f}) kernel {c5.get('kernel')}")
            e = (d.get('e2e') or {}).get('c5') or {}
            b = d.get('batched') or {}
            print(f"  C5 e2e {e.get('ms_per_step')} ms (serial {e.get('serial_ms_per_step')}); cpu {(b.get('cpu_baseline') or {}).get('ms_per_step_extrapolated')} ms (extrapolated)")
        if tm:
            print(f"  TAMOLS batched {tm.get('launch_ms'):.4f} ms frac {tm.get('frac'):.4f}")
        for k, v in (d.get('configs') or {}).items():
            print(f"  {k}: device p50 {v['device_p50_ms']:.4f} ms, e2e p50 {v['e2e_p50_ms']:.4f} ms, kernel {v['kernel']}, "
                  f"cpu p50 {(v.get('cpu_baseline') or {}).get('p50_ms')}, tick {v.get('control_tick_e2e_p50_ms')}")
        for k, v in (d.get('parity') or {}).items():
            for r in v['iterations']:
                print(f"  parity {k} it{r['iteration']}: cost rel {r['cost_max_rel_err']:.2e}, best abs {r['best_max_abs_err']:.2e}, "
                      f"tol {r['tolerance_applied']:.2e}, plain 1e-3 held {r['plain_1e-3_held']}, delta {r['delta']}, live {r['live_samples']}")
        if d.get('cpu_baseline'):
            print(f"  cpu_baseline {d['cpu_baseline']['value']:.4g} {d['cpu_baseline']['unit']} on {d['cpu_baseline']['cores']} cores")
