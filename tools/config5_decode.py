import json, sys, torch
sys.path.insert(0, ".")
import bench
dev = torch.device("cuda", 0)
F = int(sys.argv[1]) if len(sys.argv) > 1 else 512
print(json.dumps(bench.measure_config5(dev, 0, F=F)["decode"]))
