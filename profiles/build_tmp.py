import sys
sys.path.insert(0,'.')
import cuda_knearests_b200 as K
n=100_000_000
pts=K.generate_uniform_device(n)
for _ in range(4):
    h=K.KnProblem(pts,k=20,density=1.0); t=h.timings(); h.free()
print('build_ms', round(t['build_ms'],3))
