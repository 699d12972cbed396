import os, subprocess, sys, tempfile, time
ROOT="/root/repo"
sys.path.insert(0, ROOT)
import bench
enc, gpu = ROOT+"/oracle/_ref/Thorenc", ROOT+"/integration/_build/Thordec_gpu"
with tempfile.TemporaryDirectory(dir="/dev/shm") as d:
    yuv=d+"/c.yuv"; bench.write_clip(yuv,1920,1080,168)
    for n in (60, 72, 104, 120, 136, 152, 168):
        bit=d+f"/c{n}.bit"
        subprocess.run([enc,"-if",yuv,"-of",bit,"-rf",d+"/r.yuv","-stat",d+"/s.txt","-width","1920","-height","1080","-n",str(n),"-f","60","-qp","32"]+bench.DEC_CONFIGS["RA_low_complexity"],check=True,capture_output=True)
        best=1e9
        for _ in range(3):
            t0=time.perf_counter(); r=subprocess.run([gpu,bit,d+"/o.yuv"],capture_output=True,text=True,env=dict(os.environ,THORCUDA_HOOK_TIMING="1")); best=min(best,time.perf_counter()-t0)
        lv=[l for l in r.stderr.splitlines() if "levels" in l or "intra levels" in l]
        print(n, round(best,3), lv)
