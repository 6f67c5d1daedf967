"""B200-native batched world step (hot path of shacklettbp/gpu-ecs-madrona)."""
from . import scenes  # noqa: F401
from .phys import PhysExecutor, MadronaError, load_library  # noqa: F401
