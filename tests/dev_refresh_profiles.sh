#!/bin/bash
# after tests/dev_final_profiles.sh came back: gpurun_out/* -> profiles/r02_*
cp gpurun_out/launches_bench.csv profiles/r02_launches_bench.csv
python tests/dev_launch_shares.py gpurun_out/launches_bench.csv 15 > profiles/r02_launch_shares.csv
tail -1 gpurun_out/bench_ours.json > profiles/r02_bench_ours.json
tail -1 gpurun_out/bench_ref.json > profiles/r02_bench_reference.json
python tests/dev_ncu_summary.py gpurun_out/kernels_full.ncu-rep > profiles/r02_kernels_ncu.txt
for k in render_forward render_backward tile_sort sort_last_pass emit_kernel; do python tests/dev_ncu_hotspots.py gpurun_out/kernels_full.ncu-rep $k 25 | cut -c1-220; echo; done > profiles/r02_hotspots_v4.txt
python - <<'PY'
import json,re
txt=open('profiles/r02_kernels_ncu.txt').read()
blocks=re.split(r'\n(?=\S)', txt)
names={'preprocess_kernel(':'preprocess','emit_kernel':'emit','sort_last_pass_kernel':'sort_partition','tile_sort_kernel':'tile_sort','render_forward_kernel':'render_forward','render_backward_kernel':'render_backward','backward_preprocess_kernel':'backward_preprocess','bwd_order_kernel':'bwd_order','chunk_scan_kernel':'chunk_scan'}
tr={}; wi={}
for b in blocks:
    head=b.split('\n')[0]
    key=None
    for k,v in names.items():
        if k in head and not (k=='preprocess_kernel(' and 'backward' in head): key=v
    if not key: continue
    def get(m):
        x=re.search(m+r'\s+([\d.]+)',b); return float(x.group(1)) if x else 0.0
    tr[key]=int((get('dram__bytes_read.sum')+get('dram__bytes_write.sum'))*1e6)
    wi[key]=int(get('smsp__inst_executed.sum'))
out={"_source":"ncu --set full --clock-control none, one launch each at C2 (profiles/r02_kernels_ncu.txt, end of round 2); dram__bytes_read.sum + dram__bytes_write.sum",
     "warp_instructions":dict({"_source":"ncu smsp__inst_executed.sum per launch at C2 (profiles/r02_kernels_ncu.txt)"},**wi)}
out.update(tr)
json.dump(out,open('profiles/traffic.json','w'),indent=1)
d=json.loads(open('profiles/r02_bench_ours.json').read())
r=json.loads(open('profiles/r02_bench_reference.json').read())
print('ours',round(d['value'],1),round(d['ms_per_step'],4),'e2e',round(d['e2e']['value'],1),'stock',round(d['stock_signature']['value'],1),'roof',round(d['roofline']['frac'],3),'frame',round(d['frame_roofline']['frac'],3),'issue',round(d['issue_roofline']['frac'],3))
print({k:round(v*1000,1) for k,v in d['stage_ms'].items()})
print('ref',round(r['value'],1),round(r['e2e']['value'],1))
print('instr fwd/bwd',wi['render_forward'],wi['render_backward'],'dram',tr['render_forward'],tr['render_backward'])
PY
tail -11 profiles/r02_launch_shares.csv | cut -c1-120
