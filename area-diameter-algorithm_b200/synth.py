"""The reference's synthetic point generators, restated (python/generate_input_samples.py:23-48).

uniform : `count` distinct points uniform in [0, s]^2                      (rand_points_in_square)
gaussian: `count` distinct points N((c, c), std^2 I), rejected outside [0, s]^2   (norm_points_in_square)

The reference draws from numpy's legacy global stream seeded with 0 for the whole file tree and collects the
points in a Python set (so its output order is hash order).  Here each instance has its own seeded Generator
and keeps generation order; the distributions are the same, the streams are not, which is why the instances the
tests and the bench use are committed as JSON under tests/golden/.
"""
import numpy as np


def _collect(draw, count, s):
    pts, seen = [], set()
    while len(pts) < count:
        gen = draw(1000)
        for gx, gy in gen:
            p = (float(gx), float(gy))
            if 0 <= p[0] <= s and 0 <= p[1] <= s and p not in seen:
                seen.add(p)
                pts.append(p)
                if len(pts) == count:
                    break
    return pts


def uniform_points(count, seed, s=20.0):
    rng = np.random.default_rng(seed)
    return _collect(lambda k: rng.uniform(low=[0, 0], high=[s, s], size=(k, 2)), count, s)


def gaussian_points(count, std, seed, s=20.0, c=10.0):
    rng = np.random.default_rng(seed)
    return _collect(lambda k: rng.normal(loc=c, scale=std, size=(k, 2)), count, s)
