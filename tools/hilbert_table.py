"""Developer tool: derive the 24-state x 8-octant table of the 3-D Hilbert curve (build.cu kHilbertTab) from Skilling's transpose
algorithm and check it against the textbook loop."""
import itertools, sys
def spread(v, b):
    r = 0
    for i in range(b): r |= ((v >> i) & 1) << (3 * i)
    return r
def H(x, y, z, b):
    X = [x, y, z]; M = 1 << (b - 1)
    Q = M
    while Q > 1:
        P = Q - 1
        for i in range(3):
            if X[i] & Q: X[0] ^= P
            else:
                t = (X[0] ^ X[i]) & P; X[0] ^= t; X[i] ^= t
        Q >>= 1
    for i in range(1, 3): X[i] ^= X[i - 1]
    t = 0; Q = M
    while Q > 1:
        if X[2] & Q: t ^= Q - 1
        Q >>= 1
    for i in range(3): X[i] ^= t
    return spread(X[2], b) | spread(X[1], b) << 1 | spread(X[0], b) << 2
# prefix consistency
import random
random.seed(1)
for _ in range(2000):
    b = random.randint(2, 10); x, y, z = (random.randrange(1 << b) for _ in range(3))
    assert H(x, y, z, b) >> 3 == H(x >> 1, y >> 1, z >> 1, b - 1), "not prefix consistent"
print("prefix consistent")
def coords(path):   # path: list of octants o = x<<2|y<<1|z (msb first)
    x = y = z = 0
    for o in path:
        x = x << 1 | (o >> 2 & 1); y = y << 1 | (o >> 1 & 1); z = z << 1 | (o & 1)
    return x, y, z
def sig(path):
    # digits of the next two levels as a function of the next two octants
    b = len(path) + 2
    out = []
    for o1 in range(8):
        for o2 in range(8):
            x, y, z = coords(path + [o1, o2])
            out.append(H(x, y, z, b) & 63)
    return tuple(out)
states = {}; order = []; trans = {}
root = sig([])
states[root] = 0; order.append([]); queue = [[]]
while queue:
    path = queue.pop(0); s = states[sig(path)]
    for o in range(8):
        b = len(path) + 1
        x, y, z = coords(path + [o])
        digit = H(x, y, z, b) & 7
        cs = sig(path + [o])
        if cs not in states:
            states[cs] = len(states); queue.append(path + [o])
        trans[(s, o)] = (digit, states[cs])
print("states:", len(states))
ns = len(states)
tab = [trans[(s, o)][0] | trans[(s, o)][1] << 3 for s in range(ns) for o in range(8)]
print(tab)
# verify table against H for random 13-bit coords
for _ in range(20000):
    NB = 16
    x, y, z = (random.randrange(1 << NB) for _ in range(3))
    s = 0; key = 0
    for l in range(NB - 1, -1, -1):
        o = (x >> l & 1) << 2 | (y >> l & 1) << 1 | (z >> l & 1)
        e = tab[s * 8 + o]; key = key << 3 | (e & 7); s = e >> 3
    assert key == H(x, y, z, NB)
print("table ok")
