import json, sys, signal
signal.signal(signal.SIGPIPE, signal.SIG_DFL)
for f in sys.argv[1:]:
    d = json.loads([l for l in open(f) if l.startswith('{')][-1])
    print(f, f"value={d['value']:.0f} ms/step={d['ms_per_step']:.3f} e2e={d['e2e']['value']:.0f} dom={d['roofline']['kernel']} frac={d['roofline']['frac']:.3f} whole_view_frac={d['roofline']['whole_view']['frac']:.3f} latency_ms={d['config'].get('single_view_latency_ms', 0):.4f} launches/view={d['gpu_launches'] / max(1, d['steps'] * d['config'].get('views_per_gpu_per_step', 1)):.0f}")
    for k, v in d['kernels'].items():
        print(f"  {k:24s} n/view={v['launches_per_view']:.0f} avg_ms={v['avg_ms']:.4f} GB/s={v['gbs']:.0f} share={v['share_of_kernel_time']:.3f}")
