"""Experiment: bench.py at N=1 in a process that first sets up pieces of what a torchrun rank has
(EXP=torch: torch CUDA context; pg: 1-rank NCCL process group with device_id; pglazy: process group
without a communicator; libcomm: only the library's own 1-rank NCCL communicator)."""
import ctypes, runpy, sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
exp = os.environ.get('EXP', 'torch')
if exp in ('torch', 'pg', 'pglazy'):
    import torch
    torch.cuda.init(); torch.cuda.set_device(0); x = torch.ones(4, device='cuda'); torch.cuda.synchronize()
if exp in ('pg', 'pglazy'):
    import torch.distributed as dist
    os.environ.setdefault('MASTER_ADDR', '127.0.0.1'); os.environ.setdefault('MASTER_PORT', '29577')
    kw = {'device_id': torch.device('cuda', 0)} if exp == 'pg' else {}
    dist.init_process_group(backend='nccl', rank=0, world_size=1, **kw)
    if exp == 'pg':
        dist.all_reduce(x); torch.cuda.synchronize()
if exp == 'libcomm':
    from minpy_b200._lib import get_lib
    from minpy_b200.cuda_array import ensure_device
    ensure_device(0)
    lib = get_lib()
    uid = ctypes.create_string_buffer(128)
    lib.mc_comm_unique_id(uid)
    lib.mc_comm_init(0, 1, uid)
sys.argv = ['bench.py'] + sys.argv[1:]
runpy.run_path(os.path.join(ROOT, 'bench.py'), run_name='__main__')
