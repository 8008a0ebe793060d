"""The five synthetic configurations of BASELINE.json / SURVEY.md §8(d) and their plugin sets."""
from .api import Synth, plugin_cfg

# seeds: nodes 0xF6500000 + cfg, pods 0xF6510000 + cfg
CONFIGS = {
    # C1: 1k nodes × 8 GPUs, 10k fractional-GPU pods; FlexGPU + Fit filter, FlexGPU + Allocatable(Least) scores
    1: dict(n_nodes=1000, n_pending=10000, pm_fractional=1000, pm_whole=0, pm_nogpu=0, pm_degenerate=0),
    # C2: 5k nodes, 100k mixed pods; FlexGPU + Fit filter, FlexGPU + LeastAllocated scores
    2: dict(n_nodes=5000, n_pending=100000, pm_fractional=600, pm_whole=300, pm_nogpu=50, pm_degenerate=50),
    # C3: coscheduling — 2k PodGroups (minMember 4..64) over 10k nodes
    3: dict(n_nodes=10000, n_pending=70000, n_groups=2000, n_namespaces=50, pm_neg_left_nodes=10),
    # C4: capacity scheduling — 500 namespaces over 20k nodes, 100k pods
    4: dict(n_nodes=20000, n_pending=100000, n_namespaces=500, pm_pods_without_eq=50),
    # C5: 50k nodes, 1M pods, node-sharded
    5: dict(n_nodes=50000, n_pending=1000000, pm_fractional=600, pm_whole=300, pm_nogpu=50, pm_degenerate=50),
}


def config_synth(cfg_id, n_nodes=None, n_pending=None, **over):
    kw = dict(CONFIGS[cfg_id])
    kw["seed_nodes"] = 0xF6500000 + cfg_id
    kw["seed_pods"] = 0xF6510000 + cfg_id
    if n_nodes is not None:
        kw["n_nodes"] = n_nodes
    if n_pending is not None:
        kw["n_pending"] = n_pending
    kw.update(over)
    return Synth(**kw)


def config_plugins(cfg_id):
    if cfg_id == 1:
        return plugin_cfg(filter_plugins=("NodeResourcesFit", "FlexGPU"),
                          score_plugins=(("FlexGPU", 1), ("NodeResourcesAllocatable", 1)), alloc_mode="Least")
    return plugin_cfg(filter_plugins=("NodeResourcesFit", "FlexGPU"), score_plugins=(("FlexGPU", 1), ("NodeResourcesFit", 1)),
                      fit_strategy="LeastAllocated", fit_resources=(("cpu", 1), ("memory", 1)))
