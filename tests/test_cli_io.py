"""T0: CLI grammar / usage text / output name / .raw round trips incl. >2^31-byte offsets."""
import os

import numpy as np
import pytest

from b200conv.utils import output_name, read_block, read_raw, write_block, write_raw
from b200conv.utils.cli import parse_cli, usage_text


def test_cli_positionals_like_reference():
    d = parse_cli(["prog", "waterfall_1920_2520.raw", "1920", "2520", "40", "rgb"])
    assert (d["image"], d["width"], d["height"], d["repetitions"], d["channels"]) == \
        ("waterfall_1920_2520.raw", 1920, 2520, 40, 3)
    assert parse_cli(["prog", "x.raw", "10", "20", "3", "grey"])["channels"] == 1
    # atoi semantics, no range checks (mpi_convolution.c:332-334)
    assert parse_cli(["prog", "x.raw", "12abc", "-5", "zz", "grey"])["width"] == 12


@pytest.mark.parametrize("argv", [
    ["prog"], ["prog", "x.raw", "1", "2", "3"], ["prog", "x.raw", "1", "2", "3", "gray"],
    ["prog", "x.raw", "1", "2", "3", "RGB"], ["prog", "x.raw", "1", "2", "3", "grey", "stray"],
    ["prog", "x.raw", "1", "2", "3", "grey", "--boundary", "sphere"],
])
def test_cli_rejects(argv):
    assert parse_cli(argv) is None


def test_cli_extensions_and_env(monkeypatch):
    d = parse_cli(["p", "x.raw", "8", "8", "2", "grey", "--gpus", "4", "--boundary", "torus", "--tb", "8",
                   "--check-every", "20", "--filter", "box", "--kernel", "naive", "--json", "o.json", "--profile"])
    assert (d["gpus"], d["boundary"], d["tb"], d["check_every"], d["filter"], d["kernel"], d["json"], d["profile"]) \
        == (4, 1, 8, 20, 1, 1, "o.json", 1)
    monkeypatch.setenv("B200CONV_BOUNDARY", "torus")
    monkeypatch.setenv("B200CONV_TB", "2")
    d = parse_cli(["p", "x.raw", "8", "8", "2", "grey", "--tb", "4"])
    assert d["boundary"] == 1 and d["tb"] == 4  # flag beats env
    assert d["halo_depth"] == 0                 # auto
    monkeypatch.setenv("B200CONV_HALO_DEPTH", "8")
    assert parse_cli(["p", "x.raw", "8", "8", "2", "grey"])["halo_depth"] == 8
    assert parse_cli(["p", "x.raw", "8", "8", "2", "grey", "--halo-depth", "16"])["halo_depth"] == 16
    assert parse_cli(["p", "x.raw", "8", "8", "2", "grey", "--halo-depth", "-1"]) is None


def test_usage_text_matches_reference_wording():
    assert usage_text("./mpi_convolution", 0) == \
        "\nError Input!\n./mpi_convolution Image_name width height repetitions [rgb/grey].\n\n"
    assert usage_text("./cuda_convolution", 1) == \
        "Error Input!\n./cuda_convolution image_name width height repetitions [rgb/grey].\n"


def test_output_name_rule():
    assert output_name("waterfall.raw") == "blur_waterfall.raw"
    assert output_name("dir/x.raw") == "blur_dir/x.raw"  # whole argv[1], like the reference


@pytest.mark.parametrize("ch", [1, 3])
def test_raw_roundtrip_and_blocks(tmp_path, ch):
    rng = np.random.default_rng(1)
    h, w = 12, 10
    img = rng.integers(0, 256, (h, w) if ch == 1 else (h, w, ch), dtype=np.uint8)
    p = tmp_path / "a.raw"
    write_raw(p, img)
    assert os.path.getsize(p) == h * w * ch
    assert (np.fromfile(p, dtype=np.uint8) == img.reshape(-1)).all()  # headerless, interleaved, no padding
    assert (read_raw(p, w, h, ch) == img).all()
    blk = read_block(p, w, h, ch, 4, 5, 6, 5)
    assert (blk == img[4:10, 5:10]).all()
    q = tmp_path / "b.raw"
    for r0, c0 in ((0, 0), (0, 5), (6, 0), (6, 5)):  # four ranks write their blocks
        write_block(q, img[r0:r0 + 6, c0:c0 + 5], w, h, ch, r0, c0)
    assert (read_raw(q, w, h, ch) == img).all()


def test_short_file_is_an_error(tmp_path):
    p = tmp_path / "s.raw"
    p.write_bytes(b"\x01" * 10)
    from b200conv import B2CError
    with pytest.raises(B2CError, match="short read"):
        read_raw(p, 4, 4, 1)


def test_offsets_beyond_2_gib(tmp_path):
    # 65536 x 32769 grey = 2^31 + 65536 bytes; the reference's int offsets overflow here.
    w, h = 65536, 32769
    p = tmp_path / "big.raw"
    with open(p, "wb") as f:
        f.truncate(w * h)  # sparse
    row = np.arange(64, dtype=np.uint8).reshape(1, 64)
    write_block(p, row, w, h, 1, h - 1, w - 64, create=False)
    assert (read_block(p, w, h, 1, h - 1, w - 64, 1, 64) == row).all()
    with open(p, "rb") as f:
        f.seek(w * h - 64)
        assert f.read(64) == row.tobytes()
