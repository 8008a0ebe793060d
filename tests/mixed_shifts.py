"""Pod batches whose values have different granularities — so that the int32 kernels pick different shifts for each — on one
node table: a cluster in round units, batch A in whole cores / GiB, batch B in quarter cores / 256 MiB, batch C in 100 m / 3 MiB."""
import random

import builders as B

GI, MI = 1 << 30, 1 << 20


def cluster(n_nodes, seed=5):
    rng = random.Random(seed)
    nodes = []
    for i in range(n_nodes):
        n_dev = rng.choice([4, 8])
        alloc = {"cpu": rng.choice([32000, 64000, 96000]), "memory": rng.choice([128, 256, 512]) * GI, "ephemeral-storage": 1000 * GI,
                 "pods": 110, ("scalar", 0): n_dev, ("scalar", 1): n_dev * 16384}
        pods = []
        for _ in range(rng.choice([0, 1, 2, 4])):
            lim = rng.choice([None, ("gpu", 1), ("mem", 4096), ("mem", 8192)])
            pods.append(B.pod([B.ctr({"cpu": rng.choice([1000, 2000, 8000]), "memory": rng.choice([1, 4, 16]) * GI},
                                     gpu_limit=lim[1] if lim and lim[0] == "gpu" else None,
                                     mem_limit=lim[1] if lim and lim[0] == "mem" else None)],
                              index=rng.randrange(n_dev) if lim else None))
        nodes.append(B.node(alloc, pods, valid=rng.random() > 0.02))
    return nodes


def batch(kind, n_pods, seed):
    rng = random.Random(seed)
    cpu = {"A": [1000, 2000, 4000, 16000], "B": [250, 500, 1500, 30000], "C": [100, 300, 700, 50000]}[kind]
    mem = {"A": [1 * GI, 2 * GI, 8 * GI, 200 * GI], "B": [256 * MI, 768 * MI, 5 * GI, 300 * GI], "C": [3 * MI, 9 * MI, 100 * MI, 400 * GI]}[kind]
    pods = []
    for _ in range(n_pods):
        lim = rng.choice([None, None, ("gpu", 1), ("gpu", 2), ("mem", 2048), ("mem", 8192), ("mem", 16384)])
        pods.append(B.pod([B.ctr({"cpu": rng.choice(cpu), "memory": rng.choice(mem)},
                                 gpu_limit=lim[1] if lim and lim[0] == "gpu" else None, mem_limit=lim[1] if lim and lim[0] == "mem" else None)]))
    return pods
