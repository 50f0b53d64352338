"""Shared helpers of the test-suite: golden fixtures, weights, sample regeneration."""
import hashlib
import os

import numpy as np

from fairify_b200.domains import DOMAINS, domain_for_model

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")


def load_weights(name, z=None):
    z = z if z is not None else np.load(os.path.join(GOLD, "weights.npz"))
    L = int(z[f"{name}/L"])
    return [z[f"{name}/w{l}"] for l in range(L)], [z[f"{name}/b{l}"] for l in range(L)]


def reference_samples(seed, box, S):
    """The sample matrix `simluate_data` (utils/prune.py:205-222) draws after np.random.seed(seed):
    the vectorised call is stream-identical to the reference's scalar loop (asserted in make_golden.py)."""
    st = np.random.get_state()
    try:
        np.random.seed(int(seed))
        out = np.random.randint(box[:, 0], box[:, 1] + 1, size=(int(S), box.shape[0]))
    finally:
        np.random.set_state(st)
    return out.astype(np.int32)


class Golden:
    def __init__(self):
        self.weights = np.load(os.path.join(GOLD, "weights.npz"))
        self.grids = np.load(os.path.join(GOLD, "grids.npz"))
        self.prune = np.load(os.path.join(GOLD, "prune_cases.npz"))
        self.stats = np.load(os.path.join(GOLD, "np_stats.npz"))
        self.causal = np.load(os.path.join(GOLD, "causal_cases.npz"))
        self.models = [str(m) for m in self.prune["models"]]
        self.singular = np.load(os.path.join(GOLD, "singular_cases.npz"))
        self.singular_models = [str(m) for m in self.singular["models"]]

    def model_names(self):
        return sorted({k.split("/")[0] for k in self.weights.files})

    def wb(self, name):
        return load_weights(name, self.weights)

    def cases(self, name):
        """Yield dicts of one model's prune cases (with the sample matrix regenerated and verified)."""
        n = int(self.prune[f"{name}/ncases"])
        for k in range(n):
            tag = f"{name}/{k}"
            c = {f.split("/")[-1]: self.prune[f] for f in self.prune.files if f.startswith(tag + "/")}
            c["samples"] = reference_samples(c["seed"], c["box"], c["S"])
            sha = hashlib.sha256(c["samples"].tobytes()).digest()
            assert sha == c["sim_sha256"].tobytes(), "regenerated samples differ from the reference's"
            c["tag"] = tag
            yield c

    def singular_cases(self, name):
        """One model's full-Z3 sound_prune* cases (make_golden.py singular): seed = canonical partition index."""
        z = self.singular
        for k in range(int(z[f"{name}/ncases"])):
            tag = f"{name}/{k}"
            c = {f.split("/")[-1]: z[f] for f in z.files if f.startswith(tag + "/")}
            c["samples"] = reference_samples(c["index"], c["box"], c["S"])
            assert hashlib.sha256(c["samples"].tobytes()).digest() == c["sim_sha256"].tobytes()
            c["tag"] = tag
            yield c

    def domain(self, name):
        return domain_for_model(name)
