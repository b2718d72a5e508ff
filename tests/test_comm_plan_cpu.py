"""The work split of snl_render_multi (snail_b200/csrc/snl_comm.cu), host logic only (snl_comm_plan): the reference's
taskMap (src/server.cpp:238-248: every run of `world` tiles dealt to all ranks in a shuffled order) over tiles of 4 packets,
each rank's list in row-band order. No device needed."""
import numpy as np
import pytest


@pytest.mark.parametrize("world,W,H", [(1, 1920, 1080), (2, 1920, 1080), (3, 1000, 700), (4, 3840, 2160), (8, 1920, 1080), (8, 3840, 2160),
                                       (8, 40, 24), (16, 648, 360)])
def test_every_packet_goes_to_exactly_one_rank(sb, world, W, H):
    px, py = (W + 15) // 16, (H + 15) // 16
    owner = -np.ones((py, px), np.int32)
    counts, bands = [], None
    for rank in range(world):
        packets, nb, first = sb.comm_plan(world, W, H, rank)
        bands = nb if bands is None else bands
        assert nb == bands and 1 <= nb <= 8 and first[0] == 0 and (np.diff(first) >= 0).all() and first[-1] == len(packets)
        counts.append(len(packets))
        for b in range(nb):  # a band's packets lie in the band's rows, bands come in order
            rows = packets[first[b]:first[b + 1], 1] // 16
            rpb = (py + nb - 1) // nb
            assert ((rows >= b * rpb) & (rows < (b + 1) * rpb)).all()
        for x, y, w, h in packets.tolist():
            assert x % 16 == 0 and y % 16 == 0 and w == min(16, W - x) and h == min(16, H - y)
            assert owner[y // 16, x // 16] == -1
            owner[y // 16, x // 16] = rank
    assert (owner >= 0).all() and sum(counts) == px * py
    # tiles of 4 packets are dealt in shuffled groups of `world`: shares within one tile row of each other
    assert max(counts) - min(counts) <= 4 * py + 4
    # a 1080p frame on 8 GPUs gets 2 bands (1020 packets per rank), a 4K frame 8
    if (world, W) == (8, 1920):
        assert bands == 2
    if W == 3840:
        assert bands == 8


def test_the_deal_is_the_same_in_every_process(sb):
    a = sb.comm_plan(4, 1280, 720, 2)
    b = sb.comm_plan(4, 1280, 720, 2)
    assert (a[0] == b[0]).all() and a[1] == b[1]
    with pytest.raises(sb.SnlError):
        sb.comm_plan(4, 1280, 720, 4)
