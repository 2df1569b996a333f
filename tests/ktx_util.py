"""Writers for the two KTX container layouts (Khronos KTX File Format Specification 1.1 and 2.0), cube maps only:
test assets for VulkanEngine::load_cubemap (csrc/host/vk_ktx.cpp).  Independent of the decoder under test."""
import struct

import numpy as np

GL = {"unorm": (0x1401, 1, 0x1908, 0x8058, 0x1908), "srgb": (0x1401, 1, 0x1908, 0x8C43, 0x1908), "f16": (0x140B, 2, 0x1908, 0x881A, 0x1908)}
VK = {"unorm": 37, "srgb": 43, "f16": 97}


def _bytes(img, kind):
    return np.ascontiguousarray(img, dtype=np.float16 if kind == "f16" else np.uint8).tobytes()


def ktx1_cube(levels, kind, big_endian=False, stored_levels=None, key_values=b""):
    """levels[l][f] -> bytes of a KTX 1.1 file.  stored_levels overrides numberOfMipmapLevels (0 = "generate")."""
    e = ">" if big_endian else "<"
    gl_type, type_size, gl_format, gl_internal, gl_base = GL[kind]
    size = levels[0][0].shape[0]
    kv = key_values + b"\0" * (-len(key_values) % 4)
    out = bytes([0xAB, 0x4B, 0x54, 0x58, 0x20, 0x31, 0x31, 0xBB, 0x0D, 0x0A, 0x1A, 0x0A])
    out += struct.pack(e + "13I", 0x04030201, gl_type, type_size, gl_format, gl_internal, gl_base, size, size, 0, 0, 6,
                       len(levels) if stored_levels is None else stored_levels, len(kv))
    out += kv
    for lv in levels:
        first = _bytes(lv[0], kind)
        out += struct.pack(e + "I", len(first))
        for f in lv:
            b = _bytes(f, kind)
            if big_endian and kind == "f16":
                b = np.frombuffer(b, np.uint16).byteswap().tobytes()
            out += b + b"\0" * (-len(b) % 4)
    return out


def _zstd():
    """ZSTD_compress of the system's libzstd through ctypes (the image has the library but no Python binding); None without it."""
    import ctypes
    try:
        lib = ctypes.CDLL("libzstd.so.1")
    except OSError:
        return None
    lib.ZSTD_compressBound.restype = ctypes.c_size_t; lib.ZSTD_compressBound.argtypes = [ctypes.c_size_t]
    lib.ZSTD_compress.restype = ctypes.c_size_t
    lib.ZSTD_compress.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_char_p, ctypes.c_size_t, ctypes.c_int]
    lib.ZSTD_isError.restype = ctypes.c_uint; lib.ZSTD_isError.argtypes = [ctypes.c_size_t]

    def compress(data: bytes) -> bytes:
        cap = lib.ZSTD_compressBound(len(data))
        buf = ctypes.create_string_buffer(cap)
        n = lib.ZSTD_compress(buf, cap, data, len(data), 3)
        assert not lib.ZSTD_isError(n)
        return buf.raw[:n]
    return compress


zstd_compress = _zstd()


def ktx2_cube(levels, kind, scheme=0):
    """levels[l][f] -> bytes of a KTX 2.0 file (empty DFD / key-value blocks; mip levels stored smallest first as the
    specification requires).  scheme: supercompressionScheme — 0 none, 2 Zstandard, 3 ZLIB (each level compressed on its own); any
    other value is written into the header with the level data left as it is (for the refusal test)."""
    import zlib
    size = levels[0][0].shape[0]
    n = len(levels)
    header = bytes([0xAB, 0x4B, 0x54, 0x58, 0x20, 0x32, 0x30, 0xBB, 0x0D, 0x0A, 0x1A, 0x0A])
    header += struct.pack("<9I", VK[kind], 2 if kind == "f16" else 1, size, size, 0, 0, 6, n, scheme)
    index_off = len(header) + 4 * 4 + 2 * 8
    data_off = index_off + n * 24
    data_off += -data_off % 16
    blobs, offsets, stored = [], [0] * n, [0] * n
    cur = data_off
    for l in reversed(range(n)):
        blob = b"".join(_bytes(f, kind) for f in levels[l])
        if scheme == 3:
            blob = zlib.compress(blob, 6)
        elif scheme == 2 and zstd_compress is not None:
            blob = zstd_compress(blob)
        stored[l] = len(blob)
        cur += -cur % 16
        offsets[l] = cur
        blobs.append((cur, blob))
        cur += len(blob)
    out = header + struct.pack("<4I2Q", 0, 0, 0, 0, 0, 0)
    for l in range(n):
        ln = len(_bytes(levels[l][0], kind)) * 6
        out += struct.pack("<3Q", offsets[l], stored[l], ln)
    for off, blob in blobs:
        out += b"\0" * (off - len(out)) + blob
    return out
