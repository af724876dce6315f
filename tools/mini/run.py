import subprocess, sysconfig, os, sys, torch, faulthandler
faulthandler.enable()
from torch.utils import cpp_extension as ce
here=os.path.dirname(os.path.abspath(__file__))
inc=[f"-I{p}" for p in ce.include_paths(device_type="cuda")]+[f"-I{sysconfig.get_paths()['include']}"]
tlib=os.path.join(os.path.dirname(torch.__file__),"lib")
src=open(os.path.join(here,"mini.cpp")).read().replace("CPU, ops","CUDA, ops")
open("/tmp/mini_cuda.cpp","w").write(src)
cmd=["g++","-std=c++17","-O2","-fPIC","-shared","-D_GLIBCXX_USE_CXX11_ABI=1","-w",*inc,"/tmp/mini_cuda.cpp",f"-L{tlib}","-ltorch","-ltorch_cpu","-lc10",f"-Wl,-rpath,{tlib}","-o","/tmp/mini_cuda.so"]
subprocess.check_call(cmd)
torch.ops.load_library("/tmp/mini_cuda.so")
a=torch.zeros(4,8,device="cuda")
print(torch.ops._mini.f1(a,5,4,True).shape, flush=True)
try:
    torch.ops._mini.f1(a,5,5,True)
except RuntimeError as e:
    print("ok raised", str(e)[:80])
