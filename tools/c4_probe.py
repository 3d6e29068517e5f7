import sys, json, argparse, os
sys.path.insert(0, "/root/repo")
os.chdir("/root/repo")
import torch, numpy as np
import zvec_b200 as z
import bench, bench_extras
args = argparse.Namespace(extras_scale=1.0)
dev = torch.device("cuda", 0)
torch.backends.cuda.matmul.allow_tf32 = False
out = bench_extras.c4(args, z, torch, dev, bench.peaks(), bench.compare_topk, bench.merge_candidates, print)
for b in [b for b in ("batch1", "batch1024") if b in out]:
    print(b, round(out[b]["value"]), round(out[b]["ms_per_batch"], 3), out[b]["kernel"], round(out[b]["roofline"]["frac"], 3), out[b]["roofline"]["note"])
print(out["parity"])
