import json
import os

import numpy as np

HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    return json.load(open(os.path.join(HERE, name)))


def u64(hexes):
    return np.array([int(x, 16) for x in hexes], dtype=np.uint64)


def u8(hexstr):
    return np.frombuffer(bytes.fromhex(hexstr), dtype=np.uint8).copy()
