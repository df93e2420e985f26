"""ChaCha20 block function (RFC 8439 section 2.3) in plain Python: the checker for the device salt stream."""
import struct


def _rotl(x, n):
    return ((x << n) | (x >> (32 - n))) & 0xFFFFFFFF


def _qr(x, a, b, c, d):
    x[a] = (x[a] + x[b]) & 0xFFFFFFFF; x[d] = _rotl(x[d] ^ x[a], 16)
    x[c] = (x[c] + x[d]) & 0xFFFFFFFF; x[b] = _rotl(x[b] ^ x[c], 12)
    x[a] = (x[a] + x[b]) & 0xFFFFFFFF; x[d] = _rotl(x[d] ^ x[a], 8)
    x[c] = (x[c] + x[d]) & 0xFFFFFFFF; x[b] = _rotl(x[b] ^ x[c], 7)


def chacha20_block_words(key_words, w12, w13, w14, w15):
    """16 output words for key (8 x u32) and the last four state words."""
    s = [0x61707865, 0x3320646E, 0x79622D32, 0x6B206574] + [int(k) for k in key_words] + [w12, w13, w14, w15]
    x = list(s)
    for _ in range(10):
        _qr(x, 0, 4, 8, 12); _qr(x, 1, 5, 9, 13); _qr(x, 2, 6, 10, 14); _qr(x, 3, 7, 11, 15)
        _qr(x, 0, 5, 10, 15); _qr(x, 1, 6, 11, 12); _qr(x, 2, 7, 8, 13); _qr(x, 3, 4, 9, 14)
    return [(a + b) & 0xFFFFFFFF for a, b in zip(x, s)]


def salt_stream_raw(key_words, n):
    """The library's layout: 64-bit block counter in words 12..13, words 14..15 zero; eight LE u64 per block."""
    out = []
    t = 0
    while len(out) < n:
        w = chacha20_block_words(key_words, t & 0xFFFFFFFF, t >> 32, 0, 0)
        out += [w[2 * j] | (w[2 * j + 1] << 32) for j in range(8)]
        t += 1
    return out[:n]


RFC8439_KEY = struct.unpack("<8I", bytes(range(32)))
# RFC 8439 2.3.2: counter 1, nonce 00:00:00:09:00:00:00:4a:00:00:00:00
RFC8439_BLOCK = [0xe4e7f110, 0x15593bd1, 0x1fdd0f50, 0xc47120a3, 0xc7f4d1c7, 0x0368c033, 0x9aaa2204, 0x4e6cd4c3,
                 0x466482d2, 0x09aa9f07, 0x05d7c214, 0xa2028bd9, 0xd19c12b5, 0xb94e16de, 0xe883d0cb, 0x4e3c50a2]
