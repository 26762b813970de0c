"""Compressed-size ratio of the level-6 kernel against zlib level 6 on the size cases of tests/test_gpu_gzip.py."""
import sys, zlib, gzip, os, json
sys.path.insert(0, '/root/repo')
from devspace_b200 import synth
from devspace_b200.pack import Packer
def z6(d):
    c = zlib.compressobj(6, zlib.DEFLATED, -15); return len(c.compress(d) + c.flush())
res = {}
with Packer(level=6) as p:
    for name, data in [("js_like_text", synth.text_like(6_000_000, 21).numpy().tobytes()),
                       ("elf_like_binary", synth.binary_like(6_000_000, 22).numpy().tobytes()),
                       ("c1_quickstart", gzip.decompress(open('/root/repo/tests/golden/c1_quickstart.tgz','rb').read()))]:
        raw = p.deflate(data, level=6); assert zlib.decompress(raw, -15) == data
        res[name] = round(len(raw) / z6(data), 4)
    for nm, man in [("c2_small", synth.tree_c2(3000, 12_000_000)), ("c3_small", synth.tree_c3(200, 40_000_000)),
                    ("c4_small", synth.tree_c4(2, 6 << 20))]:
        out = p.pack(man)
        tar = gzip.decompress(out)
        res[nm] = round((len(out) - 18) / z6(tar), 4)
print(json.dumps(res))
