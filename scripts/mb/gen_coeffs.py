"""Generates the polynomial coefficients of bayesbridge_b200/csrc/bb_math.cuh (near-minimax Chebyshev fits,
50-digit arithmetic).  Prints C arrays, lowest degree first, with the fit's absolute error."""
import mpmath as mp
mp.mp.dps = 50

def fit(name, f, a, b, n):
    c, err = mp.chebyfit(f, [a, b], n, error=True)
    c = c[::-1]
    print(f"// {name}: degree {n-1} on [{float(a):.6g}, {float(b):.6g}], max abs error {float(err):.2e}")
    print(f"constexpr double {name}[{n}] = {{" + ", ".join(mp.nstr(x, 20) for x in c) + "};")

ln2h = mp.log(2) / 2
# exp(r) = 1 + r + r^2 Q(r)
fit("EXP_Q", lambda r: (mp.exp(r) - 1 - r) / r**2 if abs(r) > mp.mpf('1e-15') else mp.mpf(1) / 2 + r / 6, -ln2h, ln2h, 10)
# log(m) = 2 s + s z P(z), s = (m-1)/(m+1), z = s^2 in [0, (sqrt2-1)^4/...]; smax = (sqrt2-1)/(sqrt2+1)
smax = (mp.sqrt(2) - 1) / (mp.sqrt(2) + 1)
fit("LOG_P", lambda z: (2 * mp.atanh(mp.sqrt(z)) / mp.sqrt(z) - 2) / z if z != 0 else mp.mpf(2) / 3, 0, smax**2 * mp.mpf('1.02'), 7)
# sin(t) = t + t z S(z), z = t^2 in [0, (pi/2)^2]
fit("SIN_S", lambda z: (mp.sin(mp.sqrt(z)) / mp.sqrt(z) - 1) / z if z != 0 else -mp.mpf(1) / 6, 0, (mp.pi / 2)**2 * mp.mpf('1.01'), 8)
# cos(t) = 1 + z C(z), z = t^2 in [0, (pi/2)^2]
fit("COS_C", lambda z: (mp.cos(mp.sqrt(z)) - 1) / z if z != 0 else -mp.mpf(1) / 2, 0, (mp.pi / 2)**2 * mp.mpf('1.01'), 9)
print("pi hi/lo", mp.nstr(mp.pi, 25), float(mp.pi), mp.nstr(mp.pi - mp.mpf(float(mp.pi)), 20))
ln2 = mp.log(2)
import struct
hi = float(ln2); 
# ln2_hi with low 21 bits cleared so that k*ln2_hi is exact for |k| < 2^11
bits = struct.unpack('>Q', struct.pack('>d', hi))[0] & ~((1 << 21) - 1)
hi2 = struct.unpack('>d', struct.pack('>Q', bits))[0]
print("ln2_hi", repr(hi2), "ln2_lo", mp.nstr(ln2 - mp.mpf(hi2), 20), "log2e", mp.nstr(1 / ln2, 20))
