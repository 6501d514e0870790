"""Structure-aware mutation fuzzing of the glTF / GLB loader under ASan + UBSan (CPU only, not part of pytest).
  cd tests/fuzz && g++ -std=c++17 -g -O1 -fsanitize=address,undefined -fno-sanitize-recover=undefined -I../../include \
      gltf_harness.cpp ../../gpu_raytracer_b200/csrc/rt_gltf.cpp ../../gpu_raytracer_b200/csrc/rt_host.cpp -o /tmp/gltf_harness
  python fuzz_gltf.py SEED N        # writes the corpus under /tmp/gltf_fuzz, prints the inputs that crash
bvh_harness.cpp is the matching harness for rt_bvh_build (random / NaN / out-of-range inputs, all four strategies):
  g++ ... bvh_harness.cpp ../../gpu_raytracer_b200/csrc/rt_bvh.cpp ../../gpu_raytracer_b200/csrc/rt_host.cpp -o /tmp/bvh_harness"""
import sys, os, json, random, struct, subprocess, copy, pathlib
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, '..', '..')); sys.path.insert(0, os.path.join(HERE, '..'))
import numpy as np
import test_gltf as tg
out = pathlib.Path('/tmp/gltf_fuzz'); out.mkdir(parents=True, exist_ok=True)
seeds = []
for how in ('glb', 'base64'):
    for it in (np.uint16, np.uint32, np.uint8):
        d = out / f'seed_{how}_{it.__name__}'; d.mkdir(exist_ok=True)
        r = tg.build_kitchen_sink(d, how, it)
        p = r[0] if isinstance(r, tuple) else r
        if isinstance(p, dict): p = p.get('path')
        seeds.append(str(p))
print(seeds)
EVIL = [-1, 0, 1, 2, 3, 5, 7, 255, 65535, 65536, 2**31 - 1, 2**31, 2**32 - 1, 2**32, 2**53, 1e30, -1e30, 1.5, float('1e308'), None, "x", [], {}, True]
def walk(o, path, acc):
    if isinstance(o, dict):
        for k, v in o.items(): acc.append((o, k)); walk(v, path + [k], acc)
    elif isinstance(o, list):
        for i, v in enumerate(o): acc.append((o, i)); walk(v, path + [i], acc)
def mutate_doc(doc, rng):
    d = copy.deepcopy(doc)
    for _ in range(rng.randint(1, 4)):
        acc = []; walk(d, [], acc)
        if not acc: break
        o, k = rng.choice(acc)
        r = rng.random()
        try:
            if r < 0.6: o[k] = rng.choice(EVIL)
            elif r < 0.75:
                if isinstance(o, dict): del o[k]
                else: o.pop(k)
            elif r < 0.85 and isinstance(o[k], list): o[k] = o[k] * rng.randint(2, 50)
            else: o[k] = copy.deepcopy(rng.choice(acc)[0])
        except Exception: pass
    return d
def split_glb(b):
    jl = struct.unpack_from('<I', b, 12)[0]
    js = b[20:20 + jl]; rest = b[20 + jl:]
    return json.loads(js), rest
def make_glb(doc, rest, rng):
    js = json.dumps(doc).encode()
    while len(js) % 4: js += b' '
    total = 12 + 8 + len(js) + len(rest)
    if rng.random() < 0.1: total = rng.choice([0, 12, total - 1, total + 1000, 2**32 - 1])
    jl = len(js)
    if rng.random() < 0.1: jl = rng.choice([0, 1, jl - 1, jl + 4, 2**32 - 1, 2**31])
    return b'glTF' + struct.pack('<II', 2, total) + struct.pack('<I', jl) + b'JSON' + js + rest
rng = random.Random(int(sys.argv[1]) if len(sys.argv) > 1 else 0)
N = int(sys.argv[2]) if len(sys.argv) > 2 else 2000
files = []
for i in range(N):
    s = rng.choice(seeds); b = open(s, 'rb').read()
    r = rng.random()
    if r < 0.25:  # raw byte mutations
        bb = bytearray(b)
        for _ in range(rng.randint(1, 8)):
            if not bb: break
            j = rng.randrange(len(bb)); bb[j] = rng.randrange(256)
        if rng.random() < 0.3: bb = bb[:rng.randrange(len(bb) + 1)]
        data = bytes(bb)
    else:
        try:
            if b[:4] == b'glTF':
                doc, rest = split_glb(b)
                if rng.random() < 0.2: rest = rest[:rng.randrange(len(rest) + 1)]
                data = make_glb(mutate_doc(doc, rng), rest, rng)
            else:
                data = json.dumps(mutate_doc(json.loads(b), rng)).encode()
        except Exception as e:
            continue
    f = out / f'm{i:05d}.bin'; f.write_bytes(data); files.append(str(f))
# run in batches so that a crash pinpoints a small set
bad = 0
for k in range(0, len(files), 50):
    batch = files[k:k + 50]
    p = subprocess.run(['/tmp/gltf_harness'] + batch, capture_output=True, text=True, timeout=600)
    if p.returncode != 0:
        for f in batch:
            q = subprocess.run(['/tmp/gltf_harness', f], capture_output=True, text=True, timeout=120)
            if q.returncode != 0:
                bad += 1
                print('CRASH', f, q.returncode); print('\n'.join(q.stderr.splitlines()[:12]))
                if bad >= 5: sys.exit(1)
print('done', len(files), 'inputs, crashes', bad)
