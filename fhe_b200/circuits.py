"""Gate circuits in luxb200_circuit_batch form (BatchDAG wire ids: inputs first, then one wire per gate) and their
evaluation on plaintext bits."""


def greater_than8():
    """go/tfhe/compare.go:106-139 as a circuit: inputs a0..a7, b0..b7, res0 (16 = Enc(0)); 48 bootstrapped gates."""
    gates = []
    nxt = [17]

    def add(op, i0, i1=None):
        gates.append((op, i0, i1))
        nxt[0] += 1
        return nxt[0] - 1
    res = 16
    for i in range(7, -1, -1):
        a, b = i, 8 + i
        nb, na = add("NOT", b), add("NOT", a)
        aG, bG, eq = add("AND", a, nb), add("AND", na, b), add("XNOR", a, b)
        res = add("AND", eq, res)
        res = add("OR", aG, res)
        res = add("AND", res, add("NOT", bG))
    return gates, res


def plain_circuit(gates, bits):
    w = [bits[i] for i in range(bits.shape[0])]
    f = {"OR": lambda x, y: x | y, "AND": lambda x, y: x & y, "NOR": lambda x, y: 1 - (x | y),
         "NAND": lambda x, y: 1 - (x & y), "XOR": lambda x, y: x ^ y, "XNOR": lambda x, y: 1 - (x ^ y)}
    for op, i0, i1 in gates:
        w.append(1 - w[i0] if op == "NOT" else w[i0] if op == "BOOTSTRAP" else f[op](w[i0], w[i1]))
    return w


