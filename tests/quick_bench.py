import sys, time
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import otti_b200 as sb
assert sb.lib().sb200_init(0)==0
for s in (10,16,20):
    n=1<<s
    t0=time.time(); inst,v,i=sb.Instance.produce_synthetic_r1cs(n,n,10,0x0771); t1=time.time()
    gens=sb.NIZKGens.new(n,n,10); t2=time.time()
    for it in range(3):
        t3=time.time(); p=sb.NIZK.prove(inst,v,i,gens,sb.Transcript.new(b"nizk_example"),bytes(range(64))); t4=time.time()
        print(s,"synth %.2f gens %.2f prove %.4f"%(t1-t0,t2-t1,t4-t3), {k:round(x*1e3,3) for k,x in sb.NIZK.last_profile().items()}, flush=True)
