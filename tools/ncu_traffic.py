"""DRAM traffic of one kernel from an `ncu --set full` report -> profiles/*.json
    python tools/ncu_traffic.py gpurun_out/iou_final.ncu-rep iou_tile_kernel profiles/r01_iou_tile_kernel_traffic.json"""
import csv, io, json, subprocess, sys
rep, kname, out = sys.argv[1:4]
txt = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hdr, units = rows[0], rows[1]
col = {h: i for i, h in enumerate(hdr)}
picked = [r for r in rows[2:] if kname in r[col['Kernel Name']]]
assert picked, 'kernel not in report'


def val(r, name):
    v = float(r[col[name]].replace(',', ''))
    u = units[col[name]].lower()
    return v * {'byte': 1, 'kbyte': 1e3, 'mbyte': 1e6, 'gbyte': 1e9}.get(u, 1)


rd = [val(r, 'dram__bytes_read.sum') for r in picked]
wr = [val(r, 'dram__bytes_write.sum') for r in picked]
dur = [float(r[col['gpu__time_duration.sum']].replace(',', '')) for r in picked]
d = {'kernel': picked[0][col['Kernel Name']], 'launches_in_report': len(picked),
     'dram_bytes_read_per_launch': sum(rd) / len(rd), 'dram_bytes_write_per_launch': sum(wr) / len(wr),
     'dram_bytes_per_launch': (sum(rd) + sum(wr)) / len(rd),
     'gpu_time_duration': sum(dur) / len(dur), 'gpu_time_unit': units[col['gpu__time_duration.sum']],
     'source': rep, 'note': 'ncu --set full --clock-control none, cold cache, serialised'}
json.dump(d, open(out, 'w'), indent=1)
print(json.dumps(d))
