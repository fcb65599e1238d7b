"""GPU parity of the pre-processing row (isb_preprocess behind isb.preprocess): Pillow's 8-bit bicubic resize and the
wrappers' normalisation, bit-exact against the Pillow / reference-generated goldens and the numpy oracle."""
import hashlib

import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import preprocess_np as P
from oracle import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pre():
    from isb import preprocess
    return preprocess


def _sha(a):
    return hashlib.sha1(np.ascontiguousarray(a).tobytes()).hexdigest()


def test_small_cases_bit_exact(pre):
    g = load_golden("preprocess")
    for i in range(int(g["num_small"])):
        img, resized, net_in = g[f"s{i}_in"], g[f"s{i}_resized"], g[f"s{i}_input"]
        shape = resized.shape[:2]
        out, res = pre.image_to_input(img, shape, return_resized=True)
        assert out.shape == (1, 3) + shape and out.dtype == torch.float32 and out.is_cuda
        assert np.array_equal(res[0].cpu().numpy(), resized), f"case {i}: {img.shape} -> {resized.shape}"
        assert np.array_equal(out.cpu().numpy(), net_in)
        pil = pre.resize_image(img, (shape[1], shape[0]))
        assert pil.size == (shape[1], shape[0]) and np.array_equal(np.array(pil), resized)
        assert np.array_equal(pre.preprocess_input(resized).cpu().numpy(), net_in)


@pytest.mark.parametrize("tag", ["full544", "full550", "up"])
def test_full_size_bit_exact(pre, tag):
    g = load_golden("preprocess")
    h, w = [int(v) for v in g[f"{tag}_hw"]]
    shape = tuple(int(v) for v in g[f"{tag}_shape"])
    img = synth.picture(int(g[f"{tag}_seed"]), h, w)
    assert _sha(img) == str(g[f"{tag}_in_sha"])
    out, res = pre.image_to_input(img, shape, return_resized=True)
    res, out = res[0].cpu().numpy(), out.cpu().numpy()
    assert np.array_equal(res[::9, ::7], g[f"{tag}_resized_sub"])
    assert _sha(res) == str(g[f"{tag}_resized_sha"])
    assert np.array_equal(out[:, :, ::9, ::7], g[f"{tag}_input_sub"])
    assert _sha(out) == str(g[f"{tag}_input_sha"])


def test_batch_and_other_sizes_vs_oracle(pre):
    """A batch of pictures in one call, strong reductions (many taps), one unchanged axis, a single-pixel source and
    other mean / std values, against the numpy oracle."""
    rng = np.random.RandomState(11)
    for (h, w), (oh, ow), b in [((120, 90), (17, 23), 3), ((64, 200), (64, 31), 2), ((77, 50), (20, 50), 2), ((1, 1), (9, 5), 1),
                               ((300, 500), (31, 29), 2), ((40, 30), (128, 97), 4)]:
        imgs = rng.randint(0, 256, (b, h, w, 3), dtype=np.uint8)
        mean, std = (100.0, 110.5, 90.25), (50.0, 60.5, 70.125)
        out, res = pre.image_to_input(imgs, (oh, ow), mean=mean, std=std, return_resized=True)
        assert out.shape == (b, 3, oh, ow) and res.shape == (b, oh, ow, 3)
        for i in range(b):
            ref = P.resize_u8(imgs[i], ow, oh)
            assert np.array_equal(res[i].cpu().numpy(), ref), f"{(h, w)} -> {(oh, ow)} image {i}"
            assert np.array_equal(out[i:i + 1].cpu().numpy(), P.normalize_chw(ref, mean, std))


def test_wide_rows_and_long_columns_vs_oracle(pre):
    """Shapes that leave the common paths of the two resize kernels: more than four input pixels per output column with
    at least one full block of columns (a thread stages several pixel groups and several 16-byte chunks per row), widths
    that make 192 columns per block the better split and widths that do not, rows wider than one 592-pixel block of the
    vertical pass, more than 32 vertical taps (coefficients read from the table instead of shared memory), and a picture
    whose first row is not 16-byte aligned in the batch."""
    rng = np.random.RandomState(23)
    for (h, w), (oh, ow), b in [((24, 1920), (11, 300), 2), ((19, 1601), (10, 260), 2), ((21, 1333), (9, 193), 3),
                               ((18, 2500), (8, 1300), 1), ((700, 40), (35, 37), 2), ((26, 1999), (13, 640), 2), ((20, 1300), (9, 190), 2)]:
        imgs = rng.randint(0, 256, (b, h, w, 3), dtype=np.uint8)
        out, res = pre.image_to_input(imgs, (oh, ow), return_resized=True)
        assert out.shape == (b, 3, oh, ow) and res.shape == (b, oh, ow, 3)
        for i in range(b):
            ref = P.resize_u8(imgs[i], ow, oh)
            assert np.array_equal(res[i].cpu().numpy(), ref), f"{(h, w)} -> {(oh, ow)} image {i}"
            assert np.array_equal(out[i:i + 1].cpu().numpy(), P.image_to_input(imgs[i], (oh, ow)))


def test_pil_inputs_and_grey_pictures(pre):
    """PIL pictures go in as they are; a picture that is not RGB is converted first like utils.py:9-14 cvtColor."""
    Image = pytest.importorskip("PIL.Image")
    grey = Image.fromarray(synth.picture(9, 50, 70)[:, :, 0])
    rgb = np.array(grey.convert("RGB"))
    out, res = pre.image_to_input(grey, (24, 32), return_resized=True)
    assert np.array_equal(res[0].cpu().numpy(), P.resize_u8(rgb, 32, 24))
    assert np.array_equal(out.cpu().numpy(), P.image_to_input(rgb, (24, 32)))
    pil = Image.fromarray(synth.picture(10, 40, 60))
    assert np.array_equal(np.array(pre.resize_image(pil, (30, 20))), np.array(pil.resize((30, 20), Image.BICUBIC)))


@pytest.mark.parametrize("offset", [1, 2, 3, 5])
def test_unaligned_picture_pointer(pre, offset):
    """The staging loads are 32-bit and aligned down: a picture buffer that starts at an odd byte address (a view into a
    larger allocation) takes the guarded loads at both ends of the buffer and a non-zero funnel shift on every row."""
    rng = np.random.RandomState(40 + offset)
    b, h, w = 2, 37, 53
    imgs = rng.randint(0, 256, (b, h, w, 3), dtype=np.uint8)
    flat = torch.zeros(offset + imgs.size, dtype=torch.uint8, device="cuda")
    flat[offset:] = torch.from_numpy(imgs.reshape(-1)).cuda()
    view = flat[offset:].view(b, h, w, 3)
    assert view.data_ptr() % 4 == offset % 4 and view.is_contiguous()
    out, res = pre.image_to_input(view, (19, 23), return_resized=True)
    for i in range(b):
        ref = P.resize_u8(imgs[i], 23, 19)
        assert np.array_equal(res[i].cpu().numpy(), ref)
        assert np.array_equal(out[i:i + 1].cpu().numpy(), P.normalize_chw(ref))


def test_benchmark_size_batch_properties(pre):
    """Eight 1080 x 1920 pictures to the 550 x 550 network input in one call: every picture of the batch equals the
    same picture processed alone (pictures are independent, the intermediate image is per picture), two of them equal
    the numpy oracle bit for bit, and a second call gives the same bytes (no state carried between calls)."""
    pics = np.stack([synth.picture(300 + i, 1080, 1920) for i in range(8)])
    dev_pics = torch.from_numpy(pics).cuda()
    out, res = pre.image_to_input(dev_pics, (550, 550), return_resized=True)
    assert out.shape == (8, 3, 550, 550) and res.shape == (8, 550, 550, 3)
    for i in (0, 3, 7):
        o1, r1 = pre.image_to_input(dev_pics[i], (550, 550), return_resized=True)
        assert torch.equal(o1[0], out[i]) and torch.equal(r1[0], res[i])
    for i in (1, 6):
        ref = P.resize_u8(pics[i], 550, 550)
        assert np.array_equal(res[i].cpu().numpy(), ref)
        assert np.array_equal(out[i:i + 1].cpu().numpy(), P.normalize_chw(ref))
    out2 = pre.image_to_input(dev_pics, (550, 550))
    assert torch.equal(out, out2)


def test_value_table_follows_mean_std(pre):
    img = synth.picture(5, 40, 60)
    a = pre.image_to_input(img, (20, 30)).cpu().numpy()
    b = pre.image_to_input(img, (20, 30), mean=(0.0, 0.0, 0.0), std=(1.0, 1.0, 1.0)).cpu().numpy()
    c = pre.image_to_input(img, (20, 30)).cpu().numpy()
    assert np.array_equal(a, c) and np.array_equal(a, P.image_to_input(img, (20, 30)))
    assert np.array_equal(b[0], np.transpose(P.resize_u8(img, 30, 20), (2, 0, 1)).astype(np.float32))


def test_error_convention(pre):
    from isb._lib import IsbError
    with pytest.raises((IsbError, ValueError, TypeError)):
        pre.image_to_input(np.zeros((4, 4, 4), np.uint8), (2, 2))
    with pytest.raises(IsbError):
        pre.image_to_input(np.zeros((4, 4, 3), np.uint8), (2, 2), std=(1.0, 0.0, 1.0))
    with pytest.raises(IsbError):
        pre.image_to_input(np.zeros((2, 60000, 3), np.uint8), (2, 100))     # 600:1 reduction: row span beyond shared memory
