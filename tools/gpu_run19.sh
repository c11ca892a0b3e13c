#!/bin/bash
# round-2 GPU session 19: checked build + determinism stress after the refactors; build record (ptxas -v) on the GPU box
O=gpurun_out; mkdir -p $O
( time bash tools/checked_run.sh $O/r3b_checked.txt ) > $O/r3b_checked.log 2>&1; tail -22 $O/r3b_checked.txt
python - <<'PY' > gpurun_out/r3b_build_record.txt 2>&1
import subprocess, re, os, sys
sys.path.insert(0, '.')
import __graft_entry__ as g
print(subprocess.run(['nvidia-smi','--query-gpu=name,driver_version','--format=csv,noheader'],stdout=subprocess.PIPE,text=True).stdout.strip())
print(subprocess.run([g.NVCC,'--version'],stdout=subprocess.PIPE,text=True).stdout.strip().splitlines()[-2])
srcs=[os.path.join(g.CSRC,'cuda',s) for s in g.CUDA_SOURCES]
out=subprocess.run([g.NVCC]+g.NVCC_FLAGS+['-Xptxas','-v','-shared','-I',g.INC,'-I',os.path.join(g.CSRC,'cuda'),'-o','/tmp/libmcpt_cuda_box.so']+srcs+['-lcudart'],stdout=subprocess.PIPE,stderr=subprocess.STDOUT,text=True).stdout
print('command: nvcc '+' '.join(g.NVCC_FLAGS)+' -Xptxas -v (all of __graft_entry__.CUDA_SOURCES), rebuilt on the GPU box')
fn=None
for line in out.splitlines():
    m=re.search(r"Compiling entry function '([^']+)' for 'sm_100a'", line)
    if m: fn=m.group(1)
    m=re.search(r"Used (\d+) registers.*", line)
    if m and fn:
        spill=''
        print('%-110s %s' % (subprocess.run(['c++filt',fn],stdout=subprocess.PIPE,text=True).stdout.strip()[:110], line.split('ptxas info    : ')[-1])); fn=None
    if 'spill' in line and 'bytes spill stores' in line and not line.strip().startswith('ptxas info    : 0 bytes stack frame, 0 bytes spill stores'):
        print('    '+line.split('ptxas info    : ')[-1])
import hashlib
a=open('/tmp/libmcpt_cuda_box.so','rb').read(); b=open(os.path.join(g.PKG,'libmcpt_cuda.so'),'rb').read()
print('rebuilt library %d bytes, shipped library %d bytes' % (len(a), len(b)))
PY
tail -30 gpurun_out/r3b_build_record.txt
