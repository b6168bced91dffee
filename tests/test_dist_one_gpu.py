"""Shard parity on a ONE-GPU box: the ray-sharded mapping step (coslam_b200/dist.py) and the sharded BundleAdjuster
(coslam_b200/mapping.py, group=...) with every rank on cuda:0 and the collectives over gloo.  The kernels, the shard
arithmetic, AR-1 / AR-2 and the replicated optimiser steps are the ones the NCCL runs use (tests/test_dist_gpu.py, 2+ GPUs);
only the transport differs.  3 ranks as well: an odd world size leaves ragged shards."""
import pytest

from tests import test_dist_gpu as multi

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize('world', [2, 3])
def test_sharded_step_matches_single_gpu_ranks_on_one_device(world):
    multi.check_sharded_step('gloo', True, world=world)


def test_sharded_bundle_adjustment_matches_single_gpu_ranks_on_one_device():
    multi.check_sharded_bundle_adjustment('gloo', True, world=2)
