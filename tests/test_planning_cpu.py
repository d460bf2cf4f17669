"""Host-side planning helpers that decide which kernel form a layer gets (pure Python: checked without a GPU).

They mirror checks made again inside the native launchers (csrc/gemm_tcgen05.cuh: pixel_box, conv_smallc_impl), so a drift between
the two shows up here as a changed decision for the reference's own layer shapes.
"""
import torch

from distributed_ml_pytorch_b200.ops import kernels as K
from distributed_ml_pytorch_b200.parallel.sync_optim import nvls_pays

F32, BF16 = torch.float32, torch.bfloat16


def test_element_geometry():
    assert K.kblock(BF16) == 64 and K.kblock(F32) == 32          # one 128-byte k-block
    assert K.granule(BF16) == 8 and K.granule(F32) == 4          # one 16-byte granule
    assert K.pad_for(BF16, 10) == 16 and K.pad_for(F32, 10) == 12 and K.pad_for(F32, 256) == 256


def test_pixel_box_matches_the_native_rule():
    # 128 consecutive output pixels form a (wo, ho, n) box for power-of-two maps, not for 112 x 112 or 27 x 27
    assert K._pixel_box(128, 8, 8) and K._pixel_box(128, 16, 16) and K._pixel_box(128, 128, 5) and K._pixel_box(128, 256, 3)
    assert K._pixel_box(128, 4, 4)                                # 8 whole images per tile
    assert not K._pixel_box(128, 112, 112) and not K._pixel_box(128, 27, 27) and not K._pixel_box(128, 14, 14)
    assert not K._pixel_box(128, 8, 12)                          # 16 rows of 8 do not divide a 12-row map


def test_first_layers_of_the_reference_models_get_the_row_window_form():
    # AlexNet conv1 (example/models.py:29): 11x11 s4 p5 on 32x32 -> 8x8 outputs; padded buffer 42 x 44 (kw padded to 16)
    assert K.rowwin_geometry(32, 32, 11, 11, 4, 4, 5, 5, F32) == (42, 44)
    assert K.rowwin_geometry(32, 32, 11, 11, 4, 4, 5, 5, BF16) == (42, 44)
    # LeNet conv1 (example/models.py:8): 5x5 s1 p0 on 32x32 -> 28x28 outputs: no box, granule form instead
    assert K.rowwin_geometry(32, 32, 5, 5, 1, 1, 0, 0, F32) is None
    # ResNet stem 7x7 s2 p3 on 224 -> 112: no box; CIFAR ResNet 3x3 s1 p1 on 32 -> 32: box
    assert K.rowwin_geometry(224, 224, 7, 7, 2, 2, 3, 3, BF16) is None
    assert K.rowwin_geometry(32, 32, 3, 3, 1, 1, 1, 1, BF16) == (34, 39)     # the last window reads 8 taps from column 31


def test_split_k_plans_for_alexnet_shapes():
    # weight gradient of conv3 (192 -> 384, 3x3) at batch 64: K = 64*4*4 = 1024 pixels; 3 x 14 tiles -> a few splits fill 148 SMs
    tiles = ((384 + 127) // 128) * ((1728 + 127) // 128)
    s = K.pick_split_k(tiles, 1024, bk=64)
    assert 1 < s <= 8 and tiles * s >= 100
    # never split below two k-blocks per CTA, never when the tiles already fill the machine
    assert K.pick_split_k(4, 128, bk=64) == 1
    assert K.pick_split_k(200, 1 << 20, bk=64) == 1
    # tf32 k-blocks are half as long in elements: the same bytes per CTA give the same split
    assert K.pick_split_k(6, 4096, bk=32) == K.pick_split_k(6, 8192, bk=64)
    # forward: conv3 at batch 64 is 8 x 3 tiles with K = 1728 -> split; a large-M layer is not
    assert K.fprop_split(1024, 384, 1728, bk=64) > 1
    assert K.fprop_split(200704, 64, 576, bk=64) == 1
    assert K.fprop_split(64, 10, 256, bk=64) == 1                # the classifier: too short a K loop to split


def test_cluster_split_is_opt_in():
    old = K.set_cluster_splitk(False)
    try:
        assert K.cluster_plan(1024, 384, 1728, 64) is None
    finally:
        K.set_cluster_splitk(old)


def test_nvls_is_chosen_from_the_measured_crossover():
    mb = 1 << 20
    assert not nvls_pays(2, 1 * mb) and not nvls_pays(2, 256 * mb)           # 2 GPUs: peer loads never lose
    assert nvls_pays(4, 1 * mb) and not nvls_pays(4, 64 * mb)                # 4 GPUs: only small vectors
    assert nvls_pays(8, 10 * mb) and nvls_pays(8, 256 * mb) and not nvls_pays(8, 1024 * mb)
