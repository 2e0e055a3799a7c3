import sys; sys.path.insert(0,'.')
import torch
from tools import bench_configs as bc
import bench
pk=bench.peaks()
sec=bc._secondary(pk)
for k,v in sec.items():
    if isinstance(v,dict): print(k, {a:(round(b,4) if isinstance(b,float) else b) for a,b in v.items() if a in ('ms','frac','ms_per_pair')})
