"""Sample-slice partition for multi-GPU rendering: rank r renders the global sample indices
[begin, end) of EVERY pixel; the CMJ stream is keyed by the global index (D2), so the union over ranks is exactly the
single-GPU sample set. The only exchange is one sum-reduction of the float4 accumulation buffer per frame."""


def sample_slice(spp, rank, world):
    """Balanced contiguous slice of range(spp) for `rank` of `world` (first spp % world ranks get one more)."""
    if world < 1 or not (0 <= rank < world) or spp < 0:
        raise ValueError("bad slice request")
    base, extra = divmod(spp, world)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def weak_slice(spp_per_rank, rank, world):
    """Weak scaling (bench.py): every rank renders spp_per_rank samples of a frame with spp_per_rank*world samples."""
    return spp_per_rank * world, rank * spp_per_rank, (rank + 1) * spp_per_rank
