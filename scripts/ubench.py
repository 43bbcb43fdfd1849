import sys; sys.path.insert(0,'.')
import nbody_simulation_b200 as nb
w = nb.World(100000)
for k,name in enumerate(["FFMA TFLOP/s","FFMA2 TFLOP/s","pair-force Ginter/s","pair-force packed Ginter/s"]):
    r = w.measure_fp32(k); print(name, r/1e12 if k<2 else r/1e9)
