#!/usr/bin/env python3
"""Regenerates the committed fixtures in tests/golden/ (run in the build container, where /root/reference and
oracle/_ref exist; the GPU box only reads the committed files).

  cts_uastc.json          payloads of the reference's own CTS fixtures for this path: the UASTC blocks of
                          tests/cts/clitests/input/ktx2/valid_*_2D_UASTC.ktx2 and the BC7 / ASTC blocks of
                          golden/transcode/transcode_uastc/output_*_to_{bc7,astc}.ktx2, plus the CTS encode goldens
                          (raw 8x8 inputs -> golden/create/encode_uastc/output_*.ktx2, level 2, --threads 1).
  known_answer.json       SURVEY 8(c) known-answer vector (one block, levels 0-4, + BC7 + ASTC).
  transcode_golden.npz    ~30k UASTC blocks (reference-encoded adversarial + image blocks at levels 0-3, random
                          payloads behind every mode prefix, invalid blocks) with the reference's BC7 / ASTC output.
  encode_golden.npz       source 4x4 blocks + the reference encoder's output at levels 0-3 (small set).
  hinted_golden.npz       (`make_golden.py hinted`) reference output of the hint-driven targets ETC1_RGB, ETC2_RGBA, BC1_RGB,
                          BC3_RGBA, BC4_R, BC5_RG, ETC2_EAC_R11 / RG11 (+ high-quality BC1 / BC3 / R11 / RG11, BC4 / R11 from alpha) for every second block of
                          transcode_golden.npz that the reference's ETC2 path accepts (`index` = positions in its `uastc`).
"""
import json, os, struct, sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import refutil as R  # noqa: E402

CTS = "/root/reference/tests/cts/clitests"


def ktx2_levels(path):
    d = open(path, "rb").read()
    assert d[:12] == b"\xabKTX 20\xbb\r\n\x1a\n", path
    vkfmt, tsize, w, h, depth, layers, faces, levels, scheme = struct.unpack_from("<9I", d, 12)
    out = []
    for i in range(max(1, levels)):
        off, ln, uln = struct.unpack_from("<3Q", d, 80 + 24 * i)
        out.append(d[off:off + ln])
    return dict(vkFormat=vkfmt, width=w, height=h, levels=out, scheme=scheme)


def main():
    cts = {"transcode": {}, "encode": {}}
    for sub in ["R8_UNORM", "R8_SRGB", "R8G8_UNORM", "R8G8_SRGB", "R8G8B8_UNORM", "R8G8B8_SRGB", "R8G8B8A8_UNORM", "R8G8B8A8_SRGB"]:
        src = ktx2_levels(f"{CTS}/input/ktx2/valid_{sub}_2D_UASTC.ktx2")
        bc7 = ktx2_levels(f"{CTS}/golden/transcode/transcode_uastc/output_{sub}_2D_UASTC_to_bc7.ktx2")
        astc = ktx2_levels(f"{CTS}/golden/transcode/transcode_uastc/output_{sub}_2D_UASTC_to_astc.ktx2")
        assert src["scheme"] == 0
        cts["transcode"][sub] = dict(width=src["width"], height=src["height"], uastc=src["levels"][0].hex(),
                                     bc7=bc7["levels"][0].hex(), astc=astc["levels"][0].hex(),
                                     # the whole 304-byte input file, for ktxTexture2_CreateFromMemory in the compiled-boundary test
                                     ktx2_file=open(f"{CTS}/input/ktx2/valid_{sub}_2D_UASTC.ktx2", "rb").read().hex())
    tj = json.load(open(f"{CTS}/tests/create/encode_uastc.json"))
    cts["encode_command"] = tj["command"]
    for sub, comps in [("R8_UNORM", 1), ("R8G8_UNORM", 2), ("R8G8B8_UNORM", 3), ("R8G8B8A8_UNORM", 4)]:
        g = ktx2_levels(f"{CTS}/golden/create/encode_uastc/output_{sub}.ktx2")
        raw = None
        for cand in [f"{CTS}/input/raw/raw_{sub}_2D_8x8.raw"]:
            if os.path.exists(cand):
                raw = open(cand, "rb").read()
        cts["encode"][sub] = dict(width=g["width"], height=g["height"], comps=comps, uastc=g["levels"][0].hex(),
                                  raw=(raw.hex() if raw else None))
    json.dump(cts, open(os.path.join(HERE, "cts_uastc.json"), "w"), indent=1)
    if "--cts-only" in sys.argv:
        return

    # known-answer vector
    i = np.arange(16)
    px = np.stack([16 * i, 255 - 9 * i, (37 * i) & 255, np.full(16, 255)], -1).astype(np.uint8).reshape(1, 64)
    ka = {"pixels": px.tobytes().hex(), "levels": {}}
    for lvl in range(5):
        u = R.ref_encode_blocks(px, lvl, 1)
        ka["levels"][str(lvl)] = dict(uastc=u.tobytes().hex(), bc7=R.ref_transcode(u, 0)[0].tobytes().hex(),
                                      astc=R.ref_transcode(u, 1)[0].tobytes().hex())
    json.dump(ka, open(os.path.join(HERE, "known_answer.json"), "w"), indent=1)

    # transcode corpus
    px = np.concatenate([R.adversarial_blocks(1, 160), R.ref_extract_blocks(R.img_mix(192, 192, 1))])
    enc = [R.ref_encode_blocks(px, lvl) for lvl in range(4)]
    valid = np.concatenate(enc + [R.random_payload_blocks(7, 256, include_invalid=False)])
    # keep only blocks the reference accepts in `valid`; the rest go to `invalid`
    rp = R.random_payload_blocks(9, 64, include_invalid=True)
    ok = np.array([R.ref_transcode(valid[k:k + 1], 0)[1] < 0 for k in range(len(valid))])
    invalid = np.concatenate([valid[~ok], rp[np.array([R.ref_transcode(rp[k:k + 1], 0)[1] >= 0 for k in range(len(rp))])]])
    valid = valid[ok]
    np.savez_compressed(os.path.join(HERE, "transcode_golden.npz"), uastc=valid, bc7=R.ref_transcode(valid, 0)[0],
                        astc=R.ref_transcode(valid, 1)[0], invalid=invalid, rgba=R.ref_unpack_rgba(valid)[0])
    print("transcode corpus:", valid.shape, "invalid:", invalid.shape, "modes:", np.bincount(R.block_modes(valid), minlength=19))

    # encode corpus (small; the big parity runs use oracle/_ref directly)
    pxs = np.concatenate([R.adversarial_blocks(3, 48), R.ref_extract_blocks(R.img_mix(64, 64, 5))])
    np.savez_compressed(os.path.join(HERE, "encode_golden.npz"), pixels=pxs,
                        **{f"level{l}": R.ref_encode_blocks(pxs, l) for l in range(5)})
    print("encode corpus:", pxs.shape)


def make_hinted():
    g = np.load(os.path.join(HERE, "transcode_golden.npz"))
    u = g["uastc"]
    index = np.nonzero(R.etc2_transcodable(u))[0][::2].astype(np.uint32)
    b = u[index]
    out = {"index": index}
    for name, tf in [("etc1", 0), ("etc2", 1), ("bc1", 2), ("bc3", 3), ("bc4", 4), ("bc5", 5), ("r11", 20), ("rg11", 21)]:
        out[name], ok = R.ref_transcode_hinted(b, tf)
        assert ok
    out["bc1_hq"] = R.ref_transcode_hinted(b, 2, 32)[0]
    out["bc3_hq"] = R.ref_transcode_hinted(b, 3, 32)[0]
    out["bc4_alpha"] = R.ref_transcode_hinted(b, 4, 4, True)[0]
    out["r11_hq"] = R.ref_transcode_hinted(b, 20, 32)[0]
    out["rg11_hq"] = R.ref_transcode_hinted(b, 21, 32)[0]
    out["r11_alpha"] = R.ref_transcode_hinted(b, 20, 4, True)[0]
    out["etc1_alpha"] = R.ref_transcode_hinted(b, 0, 4, True)[0]    # ETC1 from the alpha channel (XC:16560)
    np.savez_compressed(os.path.join(HERE, "hinted_golden.npz"), **out)
    print("hinted corpus:", b.shape, "modes:", np.bincount(R.block_modes(b), minlength=19))


if __name__ == "__main__":
    if sys.argv[1:] == ["hinted"]:
        make_hinted()
    else:
        main()
