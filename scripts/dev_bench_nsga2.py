import os, sys, json, types
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, bench
print(json.dumps(bench.bench_nsga2(types.SimpleNamespace(steps=30), torch.device("cuda", 0))["gpu"]))
