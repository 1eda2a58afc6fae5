"""GPU parity tests for the input step in front of the path: decoded uint8 HWC images go straight into
the stem kernels, which apply the dataset normalisation of reference datasets/cityscapes.py:122-123 while they read.

Oracle: oracle/ref_input.normalize_u8 (albumentations' float32 arithmetic restated) followed by the block / network
oracle.  fp32 path <= 1e-4 (the normalisation itself is bit-exact: same two float32 operations), bf16 path <= 2e-2.
"""
import numpy as np
import pytest
import torch

from oracle import ref_blocks as B
from oracle import ref_input, ref_nets
from tests.helpers import NETS, TOL_BF16, TOL_FP32, load_specs, rel_err
from tests.test_gpu_kernels import _randomise, _sd

pytestmark = pytest.mark.gpu


def test_normalised_image_is_bit_exact():
    from rtseg_b200 import ops
    img = ref_input.make_image_u8(1, 2, 19, 23)
    want = ref_input.normalize_u8(img)
    got = ops.normalize_u8(img.cuda())
    assert got.shape == want.shape and torch.equal(got.cpu(), want)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("pool,cout", [(True, 16), (False, 32)])
@pytest.mark.parametrize("h,w", [(37, 50), (64, 256)])
def test_stem_reads_uint8_hwc(dtype, pool, cout, h, w):
    """InitialBlock(3, 16) (ERFNet / LEDNet stem, with the max-pool concat) and ConvBNAct(3, 32, 3, 2) (Fast-SCNN /
    DABNet stem) from uint8 images: tensor-core stem for bf16, CUDA-core stem for fp32."""
    from rtseg_b200 import modules, nets, ops
    img = ref_input.make_image_u8(2, 2, h, w)
    xn = ref_input.normalize_u8(img)
    if pool:
        m = _randomise(nets.InitialBlock(3, cout, "relu"), 13).cuda()
        m.act_dtype = dtype
        with torch.no_grad():
            ref = B.initial_block(xn, _sd(m), "m.", "relu")
            out = m(img.cuda())                       # [N, H, W, 3] bytes, as decoded
    else:
        m = _randomise(modules.ConvBNAct(3, cout, 3, 2, act_type="relu"), 14).cuda()
        folded, _ = m._folded()
        with torch.no_grad():
            ref = B.conv_bn_act(xn, _sd(m), "m.", 3, 2, 1, "relu")
            out = ops.stem(folded, img.cuda().permute(0, 3, 1, 2), pool=False, out_dtype=dtype)   # channels_last view
    assert ops.tc_timeout_flag() == 0
    assert out.dtype == dtype and tuple(out.shape) == tuple(ref.shape)
    assert rel_err(out, ref) < (TOL_FP32 if dtype == torch.float32 else TOL_BF16)
    if pool and dtype == torch.float32:   # the pooled channels are the normalised samples themselves: exact
        assert torch.equal(out[:, cout - 3:].cpu(), ref[:, cout - 3:])


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_avgpool_pyramid_reads_uint8_hwc(dtype):
    """DABNet's input pyramid (reference models/dabnet.py:32-34) from the uint8 image."""
    import torch.nn.functional as F
    from rtseg_b200 import ops
    img = ref_input.make_image_u8(3, 2, 33, 47)
    ref = F.avg_pool2d(ref_input.normalize_u8(img), 3, 2, 1)
    buf = torch.zeros((2, 8, 17, 24), dtype=dtype, device="cuda").contiguous(memory_format=torch.channels_last)
    ops.avgpool3x3s2(img.cuda(), out=buf, y_coff=5)
    assert rel_err(buf[:, 5:8], ref) < (1e-6 if dtype == torch.float32 else TOL_BF16)
    assert (buf[:, :5] == 0).all()


@pytest.mark.parametrize("name", NETS)
def test_networks_from_uint8_images(name):
    """All four nets, fp32 and bf16, from uint8 images at 2 x 256 x 512 against the oracle on the normalised image;
    the class map of predict() equals the one computed from the float image the reference would have built."""
    from rtseg_b200 import nets, ops
    sd = ref_nets.make_state(load_specs()[name]["spec"], 2, True)
    m = nets.MODEL_HUB[name](num_class=19)
    m.load_state_dict(sd)
    m = m.eval().cuda()
    img = ref_input.make_image_u8(4, 2, 256, 512)
    xn = ref_input.normalize_u8(img)
    with torch.no_grad():
        ref = ref_nets.FORWARD[name](sd, xn)
        y32 = m(img.cuda())
        m.set_act_dtype(torch.bfloat16)
        y16 = m(img.cuda())
        labels_u8 = m.predict(img.cuda())
        labels_f = m.predict(xn.cuda())
        y16_f = m(xn.cuda())
    assert ops.tc_timeout_flag() == 0
    assert rel_err(y32, ref) < TOL_FP32
    # bf16: the uint8 route and the float route round the same normalised samples to bf16 -> identical results
    assert torch.equal(y16, y16_f) and torch.equal(labels_u8, labels_f)
    print(f"{name} from uint8: fp32 rel err {rel_err(y32, ref):.2e}, bf16 vs fp32 oracle {rel_err(y16, ref):.2e}")


def test_engine_takes_uint8_host_images():
    import rtseg_b200
    from rtseg_b200 import nets
    sd = ref_nets.make_state(load_specs()["erfnet"]["spec"], 8, True)
    m = nets.ERFNet(num_class=19)
    m.load_state_dict(sd)
    m = m.eval().cuda().set_act_dtype(torch.bfloat16)
    eng = rtseg_b200.InferenceEngine(m, batch=2, height=128, width=256, in_dtype=torch.uint8)
    imgs = [ref_input.make_image_u8(40 + i, 2, 128, 256).pin_memory() for i in range(3)]
    with torch.no_grad():
        want = [m.predict(ref_input.normalize_u8(i).cuda()).cpu() for i in imgs]
    s0, s1 = eng.submit(imgs[0]), eng.submit(imgs[1])
    got = [eng.collect(s0).clone(), eng.collect(s1).clone(), eng(imgs[2]).clone()]
    for g, w in zip(got, want):
        assert torch.equal(g, w)
    assert eng.h2d_bytes() == 2 * 128 * 256 * 3 and eng.d2h_bytes() == 2 * 128 * 256
