"""Experiment: bench.py at N=1 in a process that has torch's CUDA context initialised first (the only
difference a torchrun rank has before the process group exists)."""
import runpy, sys, os
import torch
torch.cuda.init(); torch.cuda.set_device(0); x = torch.ones(4, device='cuda'); torch.cuda.synchronize()
if os.environ.get('EXP_TORCH_STREAM') == '1':
    s = torch.cuda.Stream(); 
    with torch.cuda.stream(s): y = x * 2
    torch.cuda.synchronize()
sys.argv = ['bench.py'] + sys.argv[1:]
runpy.run_path(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'bench.py'), run_name='__main__')
