"""The 99-exchange median-of-25 selection network of csrc/dc_median.cuh, checked
exhaustively: by the zero-one principle a comparator network selects the median
of every input iff it does so for all 2^25 inputs of zeros and ones."""
import os
import re

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(os.path.dirname(HERE), "depth-conversions_b200", "csrc", "dc_median.cuh")


def test_median25_network_is_a_median_network():
    text = open(SRC).read()
    body = text[text.index("MEDIAN25_NETWORK_BEGIN"):text.index("MEDIAN25_NETWORK_END")]
    body = body[body.index("= {"):]
    net = [(int(a), int(b)) for a, b in re.findall(r"\{(\d+),\s*(\d+)\}", body)]
    assert len(net) == 99 and all(0 <= a < 25 and 0 <= b < 25 and a != b for a, b in net)
    for c in range(32):
        n = np.arange(c << 20, (c + 1) << 20, dtype=np.uint32)
        p = [((n >> i) & 1).astype(np.uint8) for i in range(25)]
        ones = np.zeros(n.shape, np.uint8)
        for x in p:
            ones += x
        for a, b in net:
            lo, hi = np.minimum(p[a], p[b]), np.maximum(p[a], p[b])
            p[a], p[b] = lo, hi
        assert np.array_equal(p[12], (ones >= 13).astype(np.uint8)), f"chunk {c}"
