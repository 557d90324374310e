import sys, os
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import corpus, oracle_lib
import __graft_entry__ as g
pkg=g.load_package(); ctx=pkg.Context(0); port=oracle_lib.port()
d=corpus.fixture('alice29.txt')
a=ctx.compress(d,mode='fast'); b=port.fast_compress(d)
print(len(a),len(b))
i=next((k for k in range(min(len(a),len(b))) if a[k]!=b[k]),None)
print('first diff at',i, a[i-8:i+8].hex(), b[i-8:i+8].hex())
