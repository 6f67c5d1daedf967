import sys, time
sys.path.insert(0, ".")
import numpy as np
from gpu_ecs_madrona_b200 import phys, scenes
for name, sc in (("hull_pile 4096 x 21", scenes.hull_pile(4096)), ("tumble 4096 x 25", scenes.tumble(4096))):
    ex = phys.PhysExecutor(sc)
    ex.run(20); ex.kernel_ms()
    ex.run_async(60); ex.sync()
    ms, n = ex.kernel_ms()
    p, c = ex.stats()
    print(name, "ms/step %.3f" % ms, "world-steps/s %.2f M" % (sc.num_worlds / ms / 1e3), "pairs %.1f contacts %.1f" % (p.mean(), c.mean()), ex.launch_info())
    ex.close()
