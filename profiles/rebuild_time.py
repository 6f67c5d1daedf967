import time, sys
sys.path.insert(0, "/root/repo")
from gpu_ecs_madrona_b200 import phys, scenes
sc = scenes.collisions64(16384)
ex = phys.PhysExecutor(sc)
ex.run(3)
for _ in range(3):
    ex.flag_rebuild(); ex.sync()
    t0 = time.perf_counter(); ex.run(1); t1 = time.perf_counter()
    ex.run(1); t2 = time.perf_counter()
    print("step with rebuild of 16384 worlds %.3f ms, plain step %.3f ms" % (1e3*(t1-t0), 1e3*(t2-t1)))
