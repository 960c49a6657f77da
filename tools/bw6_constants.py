"""BW6-761 field constants (reference: libff bw6_761_init.cpp:36-146).

Only the two moduli are taken from the reference; every Montgomery constant is
derived here and checked against the values the reference hard-codes.
"""
Q = int(
    "6891450384315732539396789682275657542479668912536150109513790"
    "1602096234222434917360876831832894116876408645677537866134511"
    "6175912055424775934951169912530159895160509937850885037254363"
    "1423596795951899700429969112842764913119068299")
R = int(
    "2586644260129690940106527336948935335363935127549146605398842"
    "62666720468348340822774968888139573360124440321458177")

FQ_LIMBS32 = 24
FR_LIMBS32 = 12
FQ_RBITS = 768
FR_RBITS = 384


def mont_consts(p, rbits):
    Rm = 1 << rbits
    return {
        "one": Rm % p,
        "r2": (Rm * Rm) % p,
        "r3": (Rm * Rm * Rm) % p,
        "inv32": (-pow(p, -1, 1 << 32)) % (1 << 32),
        "inv64": (-pow(p, -1, 1 << 64)) % (1 << 64),
    }


FQ = mont_consts(Q, FQ_RBITS)
FR = mont_consts(R, FR_RBITS)

# cross-check with the constants the reference hard-codes (bw6_761_init.cpp:58,102,115)
assert FQ["inv32"] == 0x8FA798DD and FQ["inv64"] == 0xA5593568FA798DD
assert FR["inv32"] == 0xFFFFFFFF and FR["inv64"] == 0x8508BFFFFFFFFFFF
assert Q.bit_length() == 761 and R.bit_length() == 377

FR_S = 46
FR_GENERATOR = 15
FR_ROOT_OF_UNITY = int(
    "32863578547254505029601261939868325669770508939375122462904"
    "745766352256812585773382134936404344547323199885654433")
assert pow(FR_ROOT_OF_UNITY, 1 << FR_S, R) == 1 and pow(FR_ROOT_OF_UNITY, 1 << (FR_S - 1), R) != 1


def limbs32(x, n):
    return [(x >> (32 * i)) & 0xFFFFFFFF for i in range(n)]


def limbs64(x, n):
    return [(x >> (64 * i)) & 0xFFFFFFFFFFFFFFFF for i in range(n)]
