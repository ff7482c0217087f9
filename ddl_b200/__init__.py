"""ddl_b200 - a Blackwell-native parameter-server training framework.

Capabilities of epikjjh/DIstributed-Deep-Learning (sync / async PS training of the
14-variable MNIST CNN; unsharded, block / round-robin and greedy size-balanced sharding;
``parameter_server.py`` / ``worker.py`` / ``single.py`` entry points and ``run.sh P W``),
redesigned for B200: PS shards are GPU-resident ranges of a flat arena in peer-mapped memory,
push / update / pull are single sm_100a kernels over NVLink peer pointers, and the model's hot
ops are tcgen05 / TMEM / TMA kernels (csrc/).
"""
__version__ = "0.1.0"

from .config import Config, VARIANTS  # noqa: F401
