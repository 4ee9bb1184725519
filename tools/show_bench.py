#!/usr/bin/env python
"""One-screen summary of a bench.py JSON line:  show_bench.py file.json"""
import json, sys
b = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(f"N={b['n_gpus']} value {b['value']:.0f} GB/s  {b['ms_per_step']:.3f} ms/step  roofline frac {b['roofline']['frac']:.3f}  sharded_parity {b.get('sharded_parity')}")
if b.get('plugin_device'): print(f"  plugin_device  {b['plugin_device']['ms']:.3f} ms  frac {b['plugin_device']['frac']:.3f}")
if b.get('plugin_device_resnet50'): print(f"  plugin_device_resnet50  {b['plugin_device_resnet50']['ms']:.3f} ms  frac {b['plugin_device_resnet50']['frac']:.3f}")
if b.get('e2e'): print(f"  e2e {b['e2e']['value']:.1f} GB/s  pinned H2D per GPU {[round(x, 1) for x in b['e2e']['pinned_h2d_memcpy_GBps_per_gpu_concurrent']]}")
if b.get('e2e_pageable'): print('  e2e_pageable', {k: (round(v['value'], 1), round(v['ms'], 1)) for k, v in b['e2e_pageable'].items()})
for k, v in (b.get('defenses') or {}).items(): print(f"  {k}: {v['ms']:.3f} ms  {v['GBps']:.0f} GB/s  frac {v['roofline_frac_hbm']:.3f}  cpu {v.get('cpu_baseline', {}).get('ms')}")
s = b.get('strong')
if s:
    c = s['c4']
    print(f"  C4 {c['ms']:.2f} ms  {c['GBps']:.0f} GB/s  gram {c['gram_ms_local']:.2f} ms  frac {c['roofline']['frac']:.3f} ({c['roofline']['achieved']:.0f} TF)  eff {c['efficiency']:.3f}  t1 {c.get('t1_ms_rank0_alone')}  allreduce {c['allreduce_8MB_f64_us']}  same_sel {c.get('selection_identical_to_unsharded')}  scatter {c.get('ingest_scatter')}")
    c = s['c5']
    print(f"  C5 fltrust {c['fltrust']['ms']:.3f} ms frac {c['fltrust']['roofline_frac_hbm']:.3f} eff {c['fltrust'].get('efficiency')}  clip {c['gradient_clipping']['ms']:.3f} ms frac {c['gradient_clipping']['roofline_frac_hbm']:.3f} eff {c['gradient_clipping'].get('efficiency')}  same_det {c.get('detections_identical_to_unsharded')}  scatter {c.get('ingest_scatter')}")
if b.get('cpu_baseline'): print('  cpu', b['cpu_baseline']['value'], b['cpu_baseline']['sample'])
print('  clocks', b.get('clocks'))
