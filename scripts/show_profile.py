import json, sys
d = json.load(open(sys.argv[1] if len(sys.argv) > 1 else 'gpurun_out/op_profile_ds2_n1.json'))
reps = d.get('reps', 3)
tot = sum(v['ms'] for v in d['classes'].values())
print("total ms/eval", tot / reps, "batch", d['batch'])
for k, v in d['classes'].items():
    print(f"{k:28s} {v['ms']/reps:9.3f} ms/eval {100*v['ms']/tot:5.1f}%  count/eval {v['count']/reps:.0f}")
print()
for k, v in sorted(d['ops'].items(), key=lambda kv: -kv[1]['ms'])[:int(sys.argv[2]) if len(sys.argv) > 2 else 24]:
    ms = v['ms'] / v['count']
    print(f"{k:32s} n/eval={v['count']/reps:3.0f} {ms:8.3f} ms  {v['flops']/v['count']/ms/1e9:8.1f} TF/s {v['bytes']/v['count']/ms/1e6:8.1f} GB/s  tot {v['ms']/reps:7.3f}")
