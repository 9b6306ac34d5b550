"""Mutation fuzz of the native JSON marshals through the Python binding: every input must be accepted or rejected with
an IgpuError, never crash.  usage: python scripts/fuzz_marshal.py [seed] [iterations]   |   --dump DIR writes the seed
files for the ASAN driver (scripts/fuzz_marshal_asan.cpp)."""
import json, random, sys, time
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from headlamp_intel_gpu_plugin_b200 import k8s, metrics, native, synth, columns as K

from tests.dirty import dirty_fleet
DUMP = sys.argv[2] if len(sys.argv) > 2 and sys.argv[1] == '--dump' else None
random.seed(int(sys.argv[1]) if len(sys.argv)>1 and not DUMP else 1)
nodes,pods=dirty_fleet(6,40,3)
nj,pj=json.dumps(nodes).encode(),json.dumps(pods).encode()
START,STEP=1_700_000_000_000,15000
mx=json.dumps({"status":"success","data":{"resultType":"matrix","result":[{"metric":{"chip":'a"b',"instance":"i"},"values":[[ (START+i*STEP)/1000, repr(i*1.5)] for i in range(30)]} for _ in range(5)]}}).encode()
if DUMP:
    os.makedirs(DUMP, exist_ok=True)
    for name, blob in (('nodes.json', nj), ('pods.json', pj), ('matrix.json', mx)):
        open(os.path.join(DUMP, name), 'wb').write(blob)
    sys.exit(0)
SPECIAL=b'"\\{}[],:0-9eE.+ \n\ttruefalsn\x00\xff\xc2\xa0'
def mutate(b):
    b=bytearray(b)
    for _ in range(random.randint(1,4)):
        k=random.random(); 
        if not b: break
        i=random.randrange(len(b))
        if k<0.3: b[i]=random.choice(SPECIAL)
        elif k<0.5: del b[i:i+random.randint(1,20)]
        elif k<0.7: b[i:i]=bytes(random.choice(SPECIAL) for _ in range(random.randint(1,6)))
        elif k<0.85: b=b[:i]
        else: b[i:i]=b[max(0,i-random.randint(1,40)):i]
    return bytes(b)
n_ok=n_err=0
t0=time.time()
N=int(sys.argv[2]) if len(sys.argv)>2 else 3000
for it in range(N):
    which=it%3
    try:
        if which==0: k8s.marshalFleetJson(nj, mutate(pj), options=K.OPT_POD_LIMITS|(K.OPT_GROUP_PODS_BY_NODE if it%2 else 0), threads=1+it%3)
        elif which==1: k8s.marshalFleetJson(mutate(nj), pj, threads=2)
        else: metrics.marshalMatrixJson(mutate(mx), START, STEP, 30, threads=1+it%2)
        n_ok+=1
    except native.IgpuError as e:
        assert e.status in (-1,-6,-4), e
        n_err+=1
print("fuzz done", N, "ok", n_ok, "rejected", n_err, round(time.time()-t0,1),"s")
