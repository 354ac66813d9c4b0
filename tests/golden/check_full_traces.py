"""Runs the CPU oracle through the reference's driver loop (main.cpp:180-209, shipped DtTol) on one of the large committed inputs
and compares EVERY line of Experiments/Results/<Name>/Ih0.txt (tests/golden/ih_traces.json) at its 6 printed digits, run length
included.  Usage: python tests/golden/check_full_traces.py Monitor2320   (about a minute on 8 cores; Monitor2160 and 3DMonitor220
run in full inside tests/test_oracle_kat.py).  Last run in the build container: Monitor2160 162/162, Monitor2320 327/327,
3DMonitor220 101/101 lines equal, same stopping step."""
import os, sys, time
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE)); sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import test_oracle_kat as K
name = sys.argv[1]
orc, spec = K.build_from_input(name, HERE, os.cpu_count() or 1)
ref = K.TRACES[name]["Ih"]
got=[orc.energy()]
t=time.time()
prev=float('inf'); stop=None
for i in range(spec["nSteps"]):
    r=orc.step(spec["AdmmIter"],1e-3); assert r["status"]==0
    got.append(r["Ih"])
    d=(r["Ih"]-prev)/spec["dt"]
    if i!=0 and abs(d)<spec["DtTol"]: stop=i; break
    prev=r["Ih"]
print(name,'steps run',len(got)-1,'ref lines',len(ref),'time',time.time()-t)
bad=[(i,g,e) for i,(g,e) in enumerate(zip(got,ref)) if K.fmt6(g)!=e]
print('mismatches',bad[:5], 'len equal', len(got)==len(ref))
