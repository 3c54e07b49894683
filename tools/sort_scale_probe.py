# scratch: onesweep throughput versus array size, and what sorting the same data as 512 independent buckets would cost
# (the hybrid MSD-then-LSD question: is the cliff at ~3 G pairs a locality / TLB effect of scattering over 90+ GB?)
import sys, time, os, numpy as np, torch
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, R)
import massivevoxelraytracing_b200 as hk
ctx = hk.Context(0)
o = ctx
def timed(fn):
    torch.cuda.synchronize(); t0 = time.perf_counter(); fn(); torch.cuda.synchronize(); return (time.perf_counter() - t0) * 1e3
for n in [int(float(x) * 1e6) for x in (sys.argv[1:] or ['200', '800', '1600', '2900'])]:
    g = torch.Generator(device='cuda'); g.manual_seed(1)
    keys = torch.randint(0, 1 << 45, (n,), dtype=torch.int64, device='cuda', generator=g)
    vals = torch.arange(n, dtype=torch.int64, device='cuda')
    k2, v2 = keys.clone(), vals.clone()
    timed(lambda: o.sort_pairs_device(k2, v2, 0, 45))        # warms the library's memory pool
    k2.copy_(keys); v2.copy_(vals)
    ms_full = timed(lambda: o.sort_pairs_device(k2, v2, 0, 45))
    ok = bool((k2[1:] >= k2[:-1]).all().item())
    del k2, v2
    B = 512
    per = n // B
    def buckets():
        for b in range(B):
            o.sort_pairs_device(keys[b * per:(b + 1) * per], vals[b * per:(b + 1) * per], 0, 36)
    timed(buckets)
    keys.copy_(torch.randint(0, 1 << 45, (n,), dtype=torch.int64, device='cuda', generator=g))
    ms_b = timed(buckets)
    print('n %5.0f M pairs: full 45-bit sort %.1f ms (%.0f GB/s per pass, sorted %s) | %d buckets x 36 bits: %.1f ms (%.0f GB/s per pass)' % (
        n / 1e6, ms_full, n * 32 * 5 / ms_full / 1e6, ok, B, ms_b, n * 32 * 4 / ms_b / 1e6), flush=True)
    del keys, vals
    torch.cuda.empty_cache()
