"""Probe: times the C3 IoU tile kernel with parts compiled out (ROTGEO_ABLATE
bit flags) to see which phase bounds it.  Builds variants under tools/_ablate/.
    build here:  python tools/ablate_iou.py build
    run on GPU:  python tools/ablate_iou.py run
"""
import ctypes, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, 'tools', '_ablate')
VARIANTS = {'full': 0, 'plain_loads_no_tma': 32, 'no_exact': 4, 'no_value_store': 16, 'zeros_streaming_stores': 64}
if os.environ.get('ABLATE_ONLY'):
    VARIANTS = {k: v for k, v in VARIANTS.items() if k in os.environ['ABLATE_ONLY'].split(',')}
if os.environ.get('ABLATE_REGS'):
    VARIANTS = {'full_r%s' % r: (0, r) for r in os.environ['ABLATE_REGS'].split(',')}
EXTRA = [a for a in sys.argv[2:]]

def build():
    os.makedirs(OUT, exist_ok=True)
    csrc = os.path.join(ROOT, 'orientedobjectdetection_b200', 'csrc')
    for name, flag in VARIANTS.items():
        extra = []
        if isinstance(flag, tuple):
            flag, regs = flag
            extra = ['-maxrregcount=%s' % regs]
        so = os.path.join(OUT, 'librotgeo_%s.so' % name)
        cmd = ['/usr/local/cuda/bin/nvcc', '-gencode', 'arch=compute_100a,code=sm_100a', '-O3', '-lineinfo', '-std=c++17',
               '-DROTGEO_ABLATE=%d' % flag, '-Xcompiler', '-fPIC', '-shared', '-ccbin', '/usr/bin/g++', '-o', so] + EXTRA + extra + \
              sorted(os.path.join(csrc, f) for f in os.listdir(csrc) if f.endswith('.cu'))
        subprocess.check_call(cmd)
    print('built', sorted(os.listdir(OUT)))

def run():
    import numpy as np, torch
    from orientedobjectdetection_b200 import synth
    N = 8192
    for extent in (4096.0, 1024.0):
        a, b = synth.le90_boxes(N, 1000, extent), synth.le90_boxes(N, 1, extent)
        ad, bd = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
        out = torch.empty((N, N), dtype=torch.float32, device='cuda')
        flush = torch.empty(512 << 20, dtype=torch.uint8, device='cuda')
        for name in VARIANTS:
            L = ctypes.CDLL(os.path.join(OUT, 'librotgeo_%s.so' % name))
            vp, i64, i32, f32, sz = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_float, ctypes.c_size_t
            L.rotgeo_iou_workspace_bytes.argtypes = [i64, i64]; L.rotgeo_iou_workspace_bytes.restype = sz
            L.rotgeo_packed_bytes.argtypes = [i64]; L.rotgeo_packed_bytes.restype = sz
            L.rotgeo_iou_pairwise.argtypes = [vp, i64, i32, vp, i64, i32, i32, f32, vp, i64, vp, sz, vp]
            L.rotgeo_iou_pairwise_packed.argtypes = [vp, i64, vp, i64, i32, vp, i64, vp]
            nb = L.rotgeo_iou_workspace_bytes(N, N)
            ws = torch.empty(nb, dtype=torch.uint8, device='cuda')
            st = torch.cuda.current_stream().cuda_stream
            assert L.rotgeo_iou_pairwise(ad.data_ptr(), N, 5, bd.data_ptr(), N, 5, 0, 1e-3, out.data_ptr(), N, ws.data_ptr(), nb, st) == 0
            pk1 = ws.data_ptr(); pk2 = pk1 + (L.rotgeo_packed_bytes(N) + 255) // 256 * 256
            ts = []
            for i in range(13):
                flush.zero_()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                assert L.rotgeo_iou_pairwise_packed(pk1, N, pk2, N, 0, out.data_ptr(), N, st) == 0
                e1.record(); e1.synchronize()
                if i >= 3: ts.append(e0.elapsed_time(e1) * 1e3)
            if name == 'full':
                want = out.clone()
            same = bool(torch.equal(out, want)) if name.startswith(('plain', 'no_prefetch')) else None
            print('extent %5.0f  %-20s same_as_full=%s  median %.1f us  best %.1f us' % (extent, name, same, float(np.median(ts)), min(ts)), flush=True)

if __name__ == '__main__':
    build() if sys.argv[1] == 'build' else run()
