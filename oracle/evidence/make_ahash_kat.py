"""Known-answer vectors for the ahash fallback hasher, produced by EXECUTING compiled upstream ahash-0.8.12 code.

The reference's hashing lives in the un-vendored crate `ahash = "0.8"` (/root/reference/ballista/core/Cargo.toml:47) and
cannot be built here (no Rust toolchain). This image, however, ships a Rust cdylib that statically links ahash-0.8.12 and
still carries its local symbols: wandb/bin/librust_parquet_ffi.so (built for generic x86_64: no `aesenc`, i.e. the
*fallback* hasher, the variant the reference's release build selects - SURVEY.md §0 fact 5). Two of its functions are
self-contained leaf functions (no calls, no globals), which this script calls through ctypes by address:

  ahash::random_state::RandomState::hash_one::<&[u8]>(&self, ptr, len) -> u64
      = AHasher::from_random_state(self); write_usize(len); write(bytes); finish()      ([u8]'s Hash impl)
  ahash::random_state::RandomState::from_keys(a: &[u64; 4], b: &[u64; 4], c: usize) -> RandomState
      = h = AHasher::from_random_state(RandomState{a}); h.write_usize(c);
        mix(l, r) = {h.clone(); write_u64(l); write_u64(r); finish()};  [mix(b0,b2), mix(b1,b3), mix(b2,b1), mix(b3,b0)]

Together they execute every piece of the hasher DataFusion's create_hashes relies on: the RandomState field mapping
(buffer = k1, pad = k0, extra_keys = [k2, k3]), `update` (write_u64 / write_usize), `write` for every length class,
`large_update`, and `finish`. What they do NOT execute is `RandomState::with_seeds` (k_i = seed_i ^ PI2[i]): that const fn
is not in this binary; the vectors below are taken with the keys that formula gives for seeds (0,0,0,0), and with random
keys.

Run in this image only (the library is found next to the installed `wandb` package; offsets are checked against the
function's first bytes): python oracle/evidence/make_ahash_kat.py  ->  tests/golden/ahash_0_8_12_kat.json
"""
import ctypes as C
import hashlib
import json
import os
import random
import sys

HASH_ONE, FROM_KEYS, CREATE_READER = 0x93AEE0, 0xB28940, 0x370B40
# first instructions of the two functions (mov r8,rdx; mov r10,[rdi]; ... / push r15; push r14; push r12; push rbx; ...)
SIG_HASH_ONE = bytes.fromhex("4989d04c8b174c8b5f104c8b4f18488b47084831d048bf2d7f954c2df45158")
SIG_FROM_KEYS = bytes.fromhex("415741564154534889c84989d14833460848bb2d7f954c2df45158")
PI2 = [0x452821E638D01377, 0xBE5466CF34E90C6C, 0xC0AC29B7C97C50DD, 0x3F84D5B5B5470917]
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def find_library():
    import importlib.util

    spec = importlib.util.find_spec("wandb")
    path = os.path.join(os.path.dirname(spec.origin), "bin", "librust_parquet_ffi.so")
    if not os.path.exists(path):
        sys.exit(f"{path} not found")
    return path


def main():
    path = find_library()
    raw = open(path, "rb").read()
    lib = C.CDLL(path)
    base = C.cast(lib.create_reader, C.c_void_p).value - CREATE_READER
    for off, sig in ((HASH_ONE, SIG_HASH_ONE), (FROM_KEYS, SIG_FROM_KEYS)):
        got = C.string_at(base + off, len(sig))
        if got != sig:
            sys.exit(f"unexpected code at {off:#x}: {got.hex()} (another build of the library?)")
    U4 = C.c_uint64 * 4
    hash_one = C.CFUNCTYPE(C.c_uint64, C.POINTER(U4), C.c_char_p, C.c_size_t)(base + HASH_ONE)
    from_keys = C.CFUNCTYPE(C.c_void_p, C.POINTER(U4), C.POINTER(U4), C.POINTER(U4), C.c_size_t)(base + FROM_KEYS)

    R = random.Random(20240917)
    key_sets = [PI2] + [[R.getrandbits(64) for _ in range(4)] for _ in range(3)]
    lengths = list(range(0, 41)) + [47, 48, 49, 63, 64, 65, 100, 255, 256, 1000]
    bytes_vectors, from_keys_vectors = [], []
    for keys in key_sets:
        ks = U4(*keys)
        for n in lengths:
            data = bytes(R.getrandbits(8) for _ in range(n))
            bytes_vectors.append({"keys": [f"{k:016x}" for k in keys], "data": data.hex(),
                                  "hash": f"{hash_one(C.byref(ks), data, n):016x}"})
        # text-like inputs as well (group keys / ids of the benchmark shapes)
        for s in (b"A", b"F", b"N", b"O", b"R", b"id001", b"id0000000042", b"1996-03-13", b"DELIVER IN PERSON"):
            bytes_vectors.append({"keys": [f"{k:016x}" for k in keys], "data": s.hex(),
                                  "hash": f"{hash_one(C.byref(ks), s, len(s)):016x}"})
        for _ in range(8):
            b = [R.choice([0, 1, 2, 0xFF, (1 << 63), (1 << 64) - 1, R.getrandbits(64), R.getrandbits(20)]) for _ in range(4)]
            c = R.choice([0, 1, 8, 0xFF, R.getrandbits(64), R.getrandbits(16)])
            out = U4()
            from_keys(C.byref(out), C.byref(ks), C.byref(U4(*b)), c)
            from_keys_vectors.append({"a": [f"{k:016x}" for k in keys], "b": [f"{x:016x}" for x in b], "c": f"{c:016x}",
                                      "out": [f"{x:016x}" for x in out]})
    doc = {"source": "wandb/bin/librust_parquet_ffi.so (statically links ahash-0.8.12, generic x86_64 = fallback hasher)",
           "sha256": hashlib.sha256(raw).hexdigest(),
           "functions": {"hash_one_bytes": f"{HASH_ONE:#x} ahash::random_state::RandomState::hash_one::<&[u8]>",
                         "from_keys": f"{FROM_KEYS:#x} ahash::random_state::RandomState::from_keys"},
           "note": "keys = RandomState {k0, k1, k2, k3} as stored; the first key set is PI2 = with_seeds(0,0,0,0)",
           "hash_one_bytes": bytes_vectors, "from_keys": from_keys_vectors}
    out_path = os.path.join(ROOT, "tests", "golden", "ahash_0_8_12_kat.json")
    with open(out_path, "w") as f:
        json.dump(doc, f, indent=0)
    print(f"{len(bytes_vectors)} hash_one vectors, {len(from_keys_vectors)} from_keys vectors -> {out_path}")


if __name__ == "__main__":
    main()
