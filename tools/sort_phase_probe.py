# scratch: where a tile of the bulk-copy onesweep pass spends its life.  Needs the profiling build of the library
#   make -C massivevoxelraytracing_b200/csrc OUT=$PWD/tools/_prof/libhako_b200_sortprof.so BUILD=$PWD/tools/_prof/build EXTRA=-DHAKO_SORT_PROFILE
# (thread 0 of every CTA adds clock64() deltas of its phases to a device array; never part of the shipped library).
import ctypes as C, os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, R)
os.environ.setdefault('HAKO_LIB_PATH', os.path.join(R, 'tools', '_prof', 'libhako_b200_sortprof.so'))
import torch
import massivevoxelraytracing_b200 as hk
from massivevoxelraytracing_b200 import _abi

raw = C.CDLL(os.environ['HAKO_LIB_PATH'])
raw.hako_debug_sort_profile.argtypes = [C.POINTER(C.c_ulonglong), C.c_int]
ctx = hk.Context(0)
n = int(float(sys.argv[1]) * 1e6) if len(sys.argv) > 1 else 72_700_000
bits = int(sys.argv[2]) if len(sys.argv) > 2 else 36
g = torch.Generator(device='cuda'); g.manual_seed(1)
keys = torch.randint(0, 1 << bits, (n,), dtype=torch.int64, device='cuda', generator=g)
vals = torch.arange(n, dtype=torch.int64, device='cuda')
NAMES = ['set-up (ticket, barriers, clear)', 'wait for the key tile', 'ranking (match + packed atomics)', 'column sums, publish aggregate',
         "thread 0's look-back", "everyone's look-back + tile bases", 'keys to sorted order in smem', 'key stores', 'wait for values + sync',
         'values to sorted order in smem', 'value stores']
for variant in [int(v) for v in os.environ.get('PROBE_VARIANTS', '1').split(',')]:
    os.environ['HAKO_SORT_VARIANT'] = str(variant)
    for rep in range(3):
        k2, v2 = keys.clone(), vals.clone()
        torch.cuda.synchronize()
        out = (C.c_ulonglong * 16)()
        raw.hako_debug_sort_profile(out, 1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); ctx.sort_pairs_device(k2, v2, 0, bits); e1.record(); torch.cuda.synchronize()
        raw.hako_debug_sort_profile(out, 0)
    ms = e0.elapsed_time(e1)
    tiles = out[13]
    tot = sum(out[i] for i in range(11)) + out[11] + out[15]
    print('variant %d: %d M pairs x %d bits, %.3f ms (profiling build), %d tiles, %.0f cycles per tile life = %.2f us at 1.965 GHz; look-back rounds of thread 0 per tile %.2f, polls that found nothing published %.2f; knobs %s'
          % (variant, n // 1000000, bits, ms, tiles, tot / tiles, tot / tiles / 1965.0, out[12] / tiles, out[14] / tiles,
             {k: v for k, v in os.environ.items() if k.startswith('HAKO_SORT')}), flush=True)
    print('   of the ranking row: keys shared -> registers %.0f, own warp ranked %.0f, barrier %.0f cycles' % (out[11] / tiles, out[15] / tiles, out[2] / tiles), flush=True)
    out[2] += out[11] + out[15]
    for i, nm in enumerate(NAMES):
        print('   %-38s %8.0f cycles  %5.1f %%' % (nm, out[i] / tiles, 100.0 * out[i] / tot), flush=True)
