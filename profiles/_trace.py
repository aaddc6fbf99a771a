import sys, os, ctypes, numpy as np, torch
ROOT=os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0]=[os.path.join(ROOT,'3dgs-avatar-release_b200'),ROOT]
import bench
from diff_gaussian_rasterization import _C
dev=torch.device('cuda:0')
g,wts,frames=bench.make_workload(50000,512,512,1)
hp=bench.HotPath(dev,g,wts,512,512); f=hp.frame_to_device(frames[0])
for _ in range(5): hp.render(f)
torch.cuda.synchronize()
L=_C.lib(); buf=(ctypes.c_longlong*4096)()
L.gs_debug_fwd_trace(buf,4096)
a=np.array(list(buf)).reshape(512,8)[:188,:5]
t0=a[:,0].min()
print('total cycles', a[:,4].max()-t0)
print('producer batch time median', np.median(a[:,1]-a[:,0]))
print('consumer wait median/mean', np.median(a[:,3]-a[:,2]), (a[:,3]-a[:,2]).mean())
print('consumer process median/mean', np.median(a[:,4]-a[:,3]), (a[:,4]-a[:,3]).mean())
print('consumer step (t4 to next t4) median', np.median(np.diff(a[:,4])))
print('producer ready before consumer wants it (t2 - t1) median', np.median(a[:,2]-a[:,1]))
for b in list(range(0,10))+list(range(100,108)): print(b, (a[b]-t0).tolist())
