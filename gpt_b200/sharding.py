"""Row sharding of one global minibatch over ranks — the same split libgptb200 applies internally
(compute_grad in csrc/api.cu): rank g owns rows [g*bs/G, (g+1)*bs/G) of the batch."""


def shard_range(rank: int, nranks: int, bs: int):
    return rank * bs // nranks, (rank + 1) * bs // nranks
