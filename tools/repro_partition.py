import sys, numpy as np, torch
sys.path.insert(0, '.')
import pyspark_tf_gke_b200 as pkg, pyspark_tf_gke_b200.synth as S
eng = pkg.Engine(0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 300_007
parts = int(sys.argv[2]) if len(sys.argv) > 2 else 1
keys, _ = S.host_column(3, 0, S.KIND_KEY64, n, modulus=10**8)
vals, _ = S.host_column(3, 1, S.KIND_REAL, n)
rec, counts, base = eng.hash_partition(torch.from_numpy(keys).to(eng.device), torch.from_numpy(vals).to(eng.device), parts)
torch.cuda.synchronize()
print(counts.cpu().tolist(), base.cpu().tolist(), rec[:2].cpu().tolist())
