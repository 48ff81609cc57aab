"""Gate netlists for the depth-batched evaluator (b200_circuit_run_*, csrc/circuit.cpp).

The circuit shapes are the ones libtfhe's own integration test evaluates one gate at a time
(libs/tfhe/src/test/test-addition-boot.cpp:30-115): `full_adder_MUX`, `full_adder` and
`comparison_MUX` over nb_bits-bit operands.  Here they are written once as netlists in single
assignment form (every wire is written by exactly one gate) so that the engine can levelise them and
run every independent gate of a depth level -- across all bit positions and all operand pairs of a
batch -- as one blind-rotation launch plus one key-switch launch.

Differences from the reference's C code, none of which change the computed function:
  * the initial carry is `bootsCONSTANT` (a noiseless trivial sample, boot-gates.cpp:185-189) where
    the test uses `bootsSymEncrypt(carry, 0)` with the secret key -- the cloud side has no secret key;
  * the `bootsCOPY(carry, carry + 1)` moves disappear (single assignment); the final
    `bootsCOPY(sum + nb_bits, carry)` is kept as a COPY gate so that outputs are distinct wires.

A netlist is an int32 array [num_gates][5] = (op, out_wire, in_a, in_b, in_c), unused inputs -1, over a
pool of `num_wires` wires; wire w of instance i of a replicated netlist is pool row i*num_wires + w.
"""
import ctypes as C

import numpy as np

from . import lib as _lib
from .engine import GateOp

_NIN = {GateOp.MUX: 3, GateOp.NOT: 1, GateOp.COPY: 1, GateOp.CONST0: 0, GateOp.CONST1: 0}


class Netlist:
    def __init__(self):
        self.gates = []
        self.num_wires = 0
        self.inputs = {}    # name -> list of wires
        self.outputs = {}   # name -> list of wires

    # ---- construction ----------------------------------------------------------------------------
    def new_wire(self):
        self.num_wires += 1
        return self.num_wires - 1

    def new_wires(self, count):
        return [self.new_wire() for _ in range(count)]

    def input(self, name, count):
        self.inputs[name] = self.new_wires(count)
        return self.inputs[name]

    def gate(self, op, a=-1, b=-1, c=-1):
        op = GateOp(op)
        out = self.new_wire()
        self.gates.append((int(op), out, a, b, c))
        return out

    def to_array(self):
        return np.asarray(self.gates, dtype=np.int32).reshape(-1, 5)

    def replicate(self, count):
        """`count` independent instances of this netlist over one wire pool (instance-major)."""
        g = self.to_array()
        out = np.tile(g, (count, 1)).reshape(count, -1, 5)
        off = (np.arange(count, dtype=np.int64) * self.num_wires).astype(np.int32)[:, None]
        for col in (1, 2, 3, 4):
            vals = out[:, :, col]
            out[:, :, col] = np.where(vals >= 0, vals + off, vals)
        # level order inside circuit.cpp does not depend on gate order, but topological order must hold:
        # instance-major keeps every instance's own order.
        return out.reshape(-1, 5), count * self.num_wires

    # ---- host-side analysis (C ABI helpers, no device needed) --------------------------------------
    def levelize(self):
        """(level of every gate, number of bootstrapped levels) as the engine will schedule them."""
        g = self.to_array()
        lv = np.zeros(len(g), dtype=np.int32)
        n = C.c_int32()
        _lib.check(_lib.load().b200_circuit_levelize(g.ctypes.data, len(g), self.num_wires, lv.ctypes.data, C.byref(n)))
        return lv, n.value

    def bootstrapped_gates(self):
        return sum(1 for g in self.gates if g[0] <= int(GateOp.MUX))

    def rotations(self):
        """Blind rotations per evaluation (a MUX is two, boot-gates.cpp:341-363)."""
        return sum(2 if g[0] == int(GateOp.MUX) else 1 for g in self.gates if g[0] <= int(GateOp.MUX))

    # ---- plaintext semantics (checks the netlist, not the cryptography) ----------------------------
    def evaluate_plain(self, values):
        """values: dict wire -> 0/1 (or numpy arrays of equal shape) for every input wire."""
        w = dict(values)
        for op, out, a, b, c in self.gates:
            op = GateOp(op)
            A = w.get(a)
            B = w.get(b)
            Cc = w.get(c)
            if op == GateOp.NAND: v = 1 - (A & B)
            elif op == GateOp.OR: v = A | B
            elif op == GateOp.AND: v = A & B
            elif op == GateOp.XOR: v = A ^ B
            elif op == GateOp.XNOR: v = 1 - (A ^ B)
            elif op == GateOp.NOR: v = 1 - (A | B)
            elif op == GateOp.ANDNY: v = (1 - A) & B
            elif op == GateOp.ANDYN: v = A & (1 - B)
            elif op == GateOp.ORNY: v = (1 - A) | B
            elif op == GateOp.ORYN: v = A | (1 - B)
            elif op == GateOp.MUX: v = np.where(np.asarray(A) > 0, B, Cc)
            elif op == GateOp.NOT: v = 1 - A
            elif op == GateOp.COPY: v = A
            elif op == GateOp.CONST0: v = 0
            else: v = 1
            w[out] = v
        return w


def full_adder(nb_bits=16):
    """sum[0..nb_bits] = x + y with XOR/AND gates (test-addition-boot.cpp:70-94): 5 bootstrapped gates per bit."""
    nl = Netlist()
    x, y = nl.input("x", nb_bits), nl.input("y", nb_bits)
    carry = nl.gate(GateOp.CONST0)
    s = []
    for i in range(nb_bits):
        t0 = nl.gate(GateOp.XOR, x[i], y[i])
        s.append(nl.gate(GateOp.XOR, t0, carry))
        t1 = nl.gate(GateOp.AND, x[i], y[i])
        t2 = nl.gate(GateOp.AND, carry, t0)
        carry = nl.gate(GateOp.XOR, t1, t2)
    s.append(nl.gate(GateOp.COPY, carry))
    nl.outputs["sum"] = s
    return nl


def full_adder_mux(nb_bits=16):
    """Same sum with the carry as MUX(x^y, carry, x&y) (test-addition-boot.cpp:30-67)."""
    nl = Netlist()
    x, y = nl.input("x", nb_bits), nl.input("y", nb_bits)
    carry = nl.gate(GateOp.CONST0)
    s = []
    for i in range(nb_bits):
        t0 = nl.gate(GateOp.XOR, x[i], y[i])
        s.append(nl.gate(GateOp.XOR, t0, carry))
        t1 = nl.gate(GateOp.AND, x[i], y[i])
        carry = nl.gate(GateOp.MUX, t0, carry, t1)
    s.append(nl.gate(GateOp.COPY, carry))
    nl.outputs["sum"] = s
    return nl


def comparison_mux(nb_bits=16):
    """comp = (x <= y): carry starts at 1, carry = MUX(x_i ^ y_i, y_i, carry) (test-addition-boot.cpp:97-115)."""
    nl = Netlist()
    x, y = nl.input("x", nb_bits), nl.input("y", nb_bits)
    carry = nl.gate(GateOp.CONST1)
    for i in range(nb_bits):
        t = nl.gate(GateOp.XOR, x[i], y[i])
        carry = nl.gate(GateOp.MUX, t, y[i], carry)
    nl.outputs["comp"] = [nl.gate(GateOp.COPY, carry)]
    return nl


def int_to_bits(values, nb_bits):
    """Little-endian bit planes [count][nb_bits] of unsigned integers (bit i of the value = wire i)."""
    v = np.asarray(values, dtype=np.int64).reshape(-1, 1)
    return ((v >> np.arange(nb_bits, dtype=np.int64)) & 1).astype(np.int32)


def bits_to_int(bits):
    b = np.asarray(bits, dtype=np.int64)
    return (b << np.arange(b.shape[-1], dtype=np.int64)).sum(axis=-1)


def run_batch(cloud_key, netlist, inputs, count, stream=None):
    """Evaluate `count` independent instances of `netlist`.

    inputs: dict name -> ciphertext array [count][len(netlist.inputs[name])][n+1] (numpy int32, or torch
    CUDA int32 for a device-resident pool).  Returns (outputs dict name -> [count][len][n+1], levels).
    """
    n1 = cloud_key.params.n + 1
    gates, num_wires = netlist.replicate(count)
    first = next(iter(inputs.values()))
    if isinstance(first, np.ndarray):
        # packed I/O: only input rows go up and output rows come back; the pool stays on the device
        base = (np.arange(count, dtype=np.int64) * netlist.num_wires)[:, None]
        in_names, out_names = list(netlist.inputs), list(netlist.outputs)
        in_wires = np.concatenate([base + np.asarray(netlist.inputs[k])[None, :] for k in in_names], axis=1)
        in_rows = np.concatenate([np.asarray(inputs[k], dtype=np.int32).reshape(count, -1, n1) for k in in_names], axis=1)
        out_wires = np.concatenate([base + np.asarray(netlist.outputs[k])[None, :] for k in out_names], axis=1)
        rows, levels = cloud_key.run_circuit_io(gates, num_wires, in_wires.astype(np.int32), in_rows.reshape(-1, n1),
                                                out_wires.astype(np.int32))
        rows = rows.reshape(count, -1, n1)
        outs, at = {}, 0
        for k in out_names:
            w = len(netlist.outputs[k])
            outs[k] = rows[:, at:at + w, :]
            at += w
        return outs, levels
    else:
        import torch
        pool = torch.zeros((count, netlist.num_wires, n1), dtype=torch.int32, device=first.device)
        for name, wires in netlist.inputs.items():
            pool[:, wires, :] = inputs[name]
        levels = cloud_key.run_circuit(gates, pool.view(num_wires, n1), stream=stream)
    return {name: pool[:, wires, :] for name, wires in netlist.outputs.items()}, levels
