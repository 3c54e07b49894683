# scratch: the bulk-copy onesweep variants (HAKO_SORT_VARIANT 1..3) against the first-generation kernel (0) on 72.7 M pairs x 36 bits
# (the bench scene's sort) -- device time by CUDA events, stable-sort correctness against torch.sort on a smaller ragged array
import sys, os, numpy as np, torch
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, R)
import massivevoxelraytracing_b200 as hk
ctx = hk.Context(0)
n = int(float(sys.argv[1]) * 1e6) if len(sys.argv) > 1 else 72_700_000
bits = int(sys.argv[2]) if len(sys.argv) > 2 else 36
g = torch.Generator(device='cuda'); g.manual_seed(1)
keys = torch.randint(0, 1 << bits, (n,), dtype=torch.int64, device='cuda', generator=g)
vals = torch.arange(n, dtype=torch.int64, device='cuda')
small_k = torch.randint(0, 1 << 20, (1_000_003,), dtype=torch.int64, device='cuda', generator=g)   # odd length, many duplicates
small_v = torch.arange(small_k.numel(), dtype=torch.int64, device='cuda')
ref_k, ref_i = torch.sort(small_k, stable=True)
for variant in [int(v) for v in os.environ.get('PROBE_VARIANTS', '0,1,2,3,4').split(',')]:
    os.environ['HAKO_SORT_VARIANT'] = str(variant)
    k, v = small_k.clone(), small_v.clone()
    ctx.sort_pairs_device(k, v, 0, 27); torch.cuda.synchronize()   # 27-bit window: three 9-bit passes
    ok = bool(torch.equal(k, ref_k) and torch.equal(v, ref_i))
    times = []
    for rep in range(5):
        k2, v2 = keys.clone(), vals.clone()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); ctx.sort_pairs_device(k2, v2, 0, bits); e1.record(); torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1))
    srt = bool((k2[1:] >= k2[:-1]).all().item())
    passes = (bits + 8) // 9
    best = min(times)
    print('variant %d: stable-sort check %s, big array sorted %s; %d M pairs x %d bits: best %.3f ms, median %.3f ms = %.0f GB/s algorithmic (8 + 32 P bytes per pair, P = %d)' % (
        variant, ok, srt, n // 1000000, bits, best, sorted(times)[2], n * (8 + 32 * passes) / best / 1e6, passes), flush=True)
