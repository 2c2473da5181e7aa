python -m pytest tests/test_gpu_sparse.py -m gpu -x -q 2>&1 | tail -3
for cfg in "0 8" "1 8" "1 16"; do set -- $cfg; echo "== L2PERSIST=$1 CTAS=$2"; OBGPU_L2_PERSIST=$1 OBGPU_PCG_CTAS=$2 python scripts/run_lattice.py --n 817 --max-iter 300 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['solve_ms_events']/d['pcg_iters']*1e3, 'us/iter', d['pcg_iters'])"; done
python -c "
import ctypes
import torch
p=torch.cuda.get_device_properties(0); print(p.L2_cache_size)
from cuda import cudart
print(cudart.cudaDeviceGetAttribute(cudart.cudaDeviceAttr.cudaDevAttrMaxPersistingL2CacheSize,0), cudart.cudaDeviceGetAttribute(cudart.cudaDeviceAttr.cudaDevAttrMaxAccessPolicyWindowSize,0))
"
