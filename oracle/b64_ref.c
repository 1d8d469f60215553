/*
 * oracle/b64_ref.c -- TEST INFRASTRUCTURE ONLY (the checker, never the product).
 *
 * CPU restatement, in plain C, of the byte transform behind the reference's two
 * `b64enc` call sites:
 *     deployment/helm/templates/aziot-edge-runtime-config-secret.yaml:6
 *     deployment/helm/templates/aziot-edge-vm-cloud-init-secret.yaml:6
 * and of its inverse, which Kubernetes applies when it materialises
 * `Secret.data.userdata` as the file read at
 *     deployment/helm/templates/_helper.tpl:64,73  (via aziot-edge-vm.yaml:25-28,49-51).
 *
 * The arithmetic lives in third-party code that is NOT under /root/reference:
 *   Helm v3 (version unpinned by the reference; Chart.yaml:1 `apiVersion: v2`)
 *   -> github.com/Masterminds/sprig/v3 `b64enc` (base64encode)
 *   -> Go standard library encoding/base64, `StdEncoding`
 *      (EncodeToString / DecodeString; the latter is what encoding/json uses for
 *      a `[]byte` field such as Secret.data).
 * What follows restates that package's published algorithm (RFC 4648 section 4
 * alphabet, '=' padding, no line wrapping on encode; on decode '\r' and '\n'
 * are skipped anywhere, padding is required, trailing bits are NOT checked
 * (non-strict mode), and the first offending byte is reported as
 * CorruptInputError(offset)).
 *
 * PARITY PIN: the reference holds no test or golden vector for either b64enc
 * site ("parity unpinned" by the reference's own artefacts). This
 * file is pinned instead against RFC 4648 section 10 vectors, coreutils
 * `base64 -w0`/`-d`, CPython's binascii, and the SHA-256 fingerprints of the
 * default cloud-init payload (tests/test_oracle_b64.py).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load this.
 */
#include <stddef.h>
#include <stdint.h>
#include <string.h>

static const char kStd[65] =
    "ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz0123456789+/";

/* Go: (*Encoding).EncodedLen with padding: (n + 2) / 3 * 4. */
size_t oracle_b64_encoded_len(size_t n) { return (n + 2) / 3 * 4; }

/* Go: (*Encoding).DecodedLen with padding: n / 4 * 3 (an upper bound). */
size_t oracle_b64_decoded_len_max(size_t n) { return n / 4 * 3; }

/*
 * Go: (*Encoding).Encode. Whole 3-byte groups become four alphabet characters
 * taken from bits 23..18, 17..12, 11..6, 5..0 of the big-endian 24-bit value;
 * a 1- or 2-byte remainder is zero-extended, emits 2 or 3 characters and is
 * padded with '=' to four. Returns the number of bytes written.
 */
size_t oracle_b64_encode(const uint8_t *src, size_t n, char *dst)
{
    size_t di = 0, si = 0;
    size_t whole = (n / 3) * 3;
    while (si < whole) {
        uint32_t val = ((uint32_t)src[si] << 16) | ((uint32_t)src[si + 1] << 8) |
                       (uint32_t)src[si + 2];
        dst[di + 0] = kStd[(val >> 18) & 0x3F];
        dst[di + 1] = kStd[(val >> 12) & 0x3F];
        dst[di + 2] = kStd[(val >> 6) & 0x3F];
        dst[di + 3] = kStd[val & 0x3F];
        si += 3;
        di += 4;
    }
    size_t remain = n - si;
    if (remain == 0)
        return di;
    uint32_t val = (uint32_t)src[si] << 16;
    if (remain == 2)
        val |= (uint32_t)src[si + 1] << 8;
    dst[di + 0] = kStd[(val >> 18) & 0x3F];
    dst[di + 1] = kStd[(val >> 12) & 0x3F];
    if (remain == 2) {
        dst[di + 2] = kStd[(val >> 6) & 0x3F];
        dst[di + 3] = '=';
    } else {
        dst[di + 2] = '=';
        dst[di + 3] = '=';
    }
    return di + 4;
}

static int dec_val(uint8_t c)
{
    if (c >= 'A' && c <= 'Z') return c - 'A';
    if (c >= 'a' && c <= 'z') return c - 'a' + 26;
    if (c >= '0' && c <= '9') return c - '0' + 52;
    if (c == '+') return 62;
    if (c == '/') return 63;
    return -1;
}

/*
 * Go: (*Encoding).decodeQuantum + (*Encoding).Decode for StdEncoding
 * (padded, non-strict). On success returns 0 and stores the decoded length in
 * *out_n. On malformed input returns 1 and stores Go's
 * CorruptInputError offset in *err_off; *out_n then holds the number of bytes
 * successfully written before the error, as Go's Decode returns it.
 */
int oracle_b64_decode(const uint8_t *src, size_t len, uint8_t *dst,
                      size_t *out_n, size_t *err_off)
{
    size_t si = 0, n = 0;
    for (;;) {
        /* decodeQuantum */
        uint8_t dbuf[4];
        int dlen = 4;
        int j = 0;
        int end = 0;
        for (j = 0; j < 4; j++) {
            if (si == len) {
                if (j == 0) { /* clean end of input */
                    *out_n = n;
                    return 0;
                }
                /* j == 1, or padding required: incomplete quantum */
                *out_n = n;
                *err_off = si - (size_t)j;
                return 1;
            }
            uint8_t in = src[si];
            si++;
            int v = dec_val(in);
            if (v >= 0) {
                dbuf[j] = (uint8_t)v;
                continue;
            }
            if (in == '\n' || in == '\r') {
                j--;
                continue;
            }
            if (in != '=') {
                *out_n = n;
                *err_off = si - 1;
                return 1;
            }
            /* padding */
            if (j == 0 || j == 1) {
                *out_n = n;
                *err_off = si - 1;
                return 1;
            }
            if (j == 2) {
                /* "==" expected; skip newlines first */
                while (si < len && (src[si] == '\n' || src[si] == '\r'))
                    si++;
                if (si == len) {
                    *out_n = n;
                    *err_off = len;
                    return 1;
                }
                if (src[si] != '=') {
                    *out_n = n;
                    *err_off = si - 1;
                    return 1;
                }
                si++;
            }
            /* skip over newlines; anything else after padding is an error */
            while (si < len && (src[si] == '\n' || src[si] == '\r'))
                si++;
            if (si < len) {
                end = 2; /* trailing garbage: CorruptInputError(si) */
            } else {
                end = 1;
            }
            dlen = j;
            break;
        }
        /* convert 4x 6-bit source bytes into 3 bytes */
        uint32_t val = 0;
        for (int k = 0; k < dlen; k++)
            val |= (uint32_t)dbuf[k] << (18 - 6 * k);
        switch (dlen) {
        case 4:
            dst[n + 2] = (uint8_t)(val);
            /* fallthrough */
        case 3:
            dst[n + 1] = (uint8_t)(val >> 8);
            /* fallthrough */
        case 2:
            dst[n + 0] = (uint8_t)(val >> 16);
        }
        n += (size_t)(dlen - 1);
        if (end == 2) {
            /* Go stores this quantum's bytes, then reports the garbage. */
            *out_n = n;
            *err_off = si;
            return 1;
        }
        if (end) {
            *out_n = n;
            return 0;
        }
    }
}

/*
 * Batched forms used by the parity tests and the CPU baseline: `n_units`
 * independent byte strings packed in one blob, unit i occupying
 * [in_off[i], in_off[i+1]) and writing at out_off[i]. Units are independent
 * (one install's Secret payload each).
 */
void oracle_b64_encode_batch(const uint8_t *in, const int64_t *in_off,
                             int64_t n_units, char *out, const int64_t *out_off)
{
    for (int64_t i = 0; i < n_units; i++)
        oracle_b64_encode(in + in_off[i], (size_t)(in_off[i + 1] - in_off[i]),
                          out + out_off[i]);
}

/* status[i] = -1 on success else Go's CorruptInputError offset; out_len[i] =
 * bytes written. */
void oracle_b64_decode_batch(const uint8_t *in, const int64_t *in_off,
                             int64_t n_units, uint8_t *out,
                             const int64_t *out_off, int64_t *out_len,
                             int64_t *status)
{
    for (int64_t i = 0; i < n_units; i++) {
        size_t n = 0, eo = 0;
        int rc = oracle_b64_decode(in + in_off[i],
                                   (size_t)(in_off[i + 1] - in_off[i]),
                                   out + out_off[i], &n, &eo);
        out_len[i] = (int64_t)n;
        status[i] = rc ? (int64_t)eo : -1;
    }
}
