"""Timeline of ONE call: when the host launched and when the GPU ran each kernel (torch profiler chrome trace).
    python tools/timeline.py [c4batch|c4image|c2]"""
import json, os, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from torch.profiler import profile, ProfilerActivity
import orientedobjectdetection_b200 as rg
from orientedobjectdetection_b200 import synth, dist as rdist
what = sys.argv[1] if len(sys.argv) > 1 else 'c4batch'
cu = lambda x: torch.from_numpy(x).cuda()
if what == 'c4batch':
    imgs = [tuple(cu(x) for x in synth.c4_rpn_image(s)) for s in range(64)]
    bat = rdist.Batched(torch.cat([b for b, _, _ in imgs]), torch.cat([s for _, s, _ in imgs]),
                        torch.cat([i for _, _, i in imgs]), [int(b.shape[0]) for b, _, _ in imgs])
    fn = lambda: rdist.nms_images(bat, 0.8, 2000)
elif what == 'c4image':
    b1, s1, i1 = [cu(x) for x in synth.c4_rpn_image(0)]
    fn = lambda: rg.batched_nms(b1, s1, i1, dict(type='nms_rotated', iou_threshold=0.8))
else:
    mb, ms = synth.c2_multiclass(0, dense=False)
    mbd, msd = cu(mb), cu(ms)
    fn = lambda: rg.multiclass_nms_rotated(mbd, msd, 0.05, dict(iou_thr=0.1), 2000)
for _ in range(5):
    fn()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    fn()
    torch.cuda.synchronize()
    fn()
    torch.cuda.synchronize()
path = os.path.join(tempfile.gettempdir(), 'tl.json')
prof.export_chrome_trace(path)
ev = [e for e in json.load(open(path))['traceEvents'] if e.get('ph') == 'X']
gpu = sorted([e for e in ev if e.get('cat') in ('kernel', 'gpu_memcpy', 'gpu_memset')], key=lambda e: e['ts'])
rt = sorted([e for e in ev if e.get('cat') == 'cuda_runtime'], key=lambda e: e['ts'])
# second call only
half = gpu[len(gpu) // 2:]
first_rt = [e for e in rt if e['ts'] > gpu[len(gpu) // 2 - 1]['ts'] + gpu[len(gpu) // 2 - 1]['dur']]
t0 = first_rt[0]['ts'] if first_rt else half[0]['ts']
print('== %s: GPU activities of one call (us after the call\'s first CUDA runtime call)' % what)
prev_end = None
for e in half:
    gap = '' if prev_end is None else ' (idle %.1f before)' % (e['ts'] - prev_end)
    print('  gpu  start %7.1f  dur %6.1f  %s%s' % (e['ts'] - t0, e['dur'], e['name'][:60], gap))
    prev_end = e['ts'] + e['dur']
print('== host: CUDA runtime calls')
for e in first_rt:
    print('  host start %7.1f  dur %6.1f  %s' % (e['ts'] - t0, e['dur'], e['name'][:40]))
