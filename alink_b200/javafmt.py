"""Java-compatible number/vector formatting used for `predictionDetailCol`.

Mirrors `Double.toString` (JLS: decimal form for 1e-3 <= |x| < 1e7, otherwise
"computerized scientific notation") over the shortest round-trip digits, and
`VectorUtil.toString(DenseVector)` = values joined by a single space
(reference: core/src/main/java/com/alibaba/alink/common/linalg/VectorUtil.java,
used by KMeansModelMapper.java:83-90).
"""
import math


def java_double_to_string(x):
    x = float(x)
    if math.isnan(x):
        return "NaN"
    if math.isinf(x):
        return "Infinity" if x > 0 else "-Infinity"
    if x == 0.0:
        return "-0.0" if math.copysign(1.0, x) < 0 else "0.0"
    sign = "-" if x < 0 else ""
    a = abs(x)
    mant, exp = repr(a), 0  # shortest round-trip digits
    # normalise python repr (which may be '1e-05', '1.5e+20', '123.45') to digits + exponent
    if "e" in mant or "E" in mant:
        m, e = mant.lower().split("e")
        exp = int(e)
    else:
        m = mant
    if "." in m:
        ip, fp = m.split(".")
    else:
        ip, fp = m, ""
    digits = (ip + fp).lstrip("0")
    # decimal exponent of the first significant digit
    if ip.strip("0") != "":
        e10 = len(ip.lstrip("0")) - 1 + exp
    else:
        e10 = -(len(fp) - len(fp.lstrip("0")) + 1) + exp
    digits = digits.rstrip("0") or "0"
    if 1e-3 <= a < 1e7:
        if e10 >= 0:
            ipart = digits[: e10 + 1].ljust(e10 + 1, "0")
            fpart = digits[e10 + 1:] or "0"
        else:
            ipart = "0"
            fpart = "0" * (-e10 - 1) + digits
        return "%s%s.%s" % (sign, ipart, fpart)
    return "%s%s.%sE%d" % (sign, digits[0], digits[1:] or "0", e10)


def java_vector_to_string(values):
    return " ".join(java_double_to_string(v) for v in values)
