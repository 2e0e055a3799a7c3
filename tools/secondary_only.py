import sys, json
sys.path.insert(0, '/root/repo')
import torch, bench
from tools import bench_configs as bc
bc.DEV = torch.device('cuda:0')
s = bc._secondary(bench.peaks())
print(json.dumps({k: (round(v['frac'], 3), round(v.get('ms', v.get('ms_per_pair', 0)), 4)) for k, v in s.items() if isinstance(v, dict)}))
