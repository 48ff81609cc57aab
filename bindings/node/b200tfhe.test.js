// b200tfhe.test.js -- node:test cases for the GPU gate path, shaped like the reference's harness
// (tests/tfhe.test.js:56-186: one `test()` per operation, assert on the decoded result).  Needs node,
// the built addon (see b200tfhe_napi.cc) and the fixtures written by bindings/node/make_fixtures.py
// (cloud.key, a.ct, b.ct, c.ct, bits.json, lwe_key.json, eoc.json).  Not runnable in the round-1 image (no node).
const test = require('node:test')
const assert = require('node:assert')
const fs = require('fs')
const path = require('path')
const b200 = require('./build/b200tfhe.node')

const fx = (name) => fs.readFileSync(path.join(__dirname, 'fixtures', name))
const OP = { NAND: 0, OR: 1, AND: 2, XOR: 3, XNOR: 4, NOR: 5, ANDNY: 6, ANDYN: 7, ORNY: 8, ORYN: 9, MUX: 10 }

function decrypt (cts, key, n) {          // bit = (b - <a, s>) > 0 on the torus, as bootsSymDecrypt
  const out = []
  for (let i = 0; i < cts.length / (n + 1); i++) {
    let acc = 0
    for (let j = 0; j < n; j++) acc = (acc + Math.imul(cts[i * (n + 1) + j], key[j])) | 0
    out.push((((cts[i * (n + 1) + n] - acc) | 0) > 0) ? 1 : 0)
  }
  return out
}

test('info reports a CUDA device', () => {
  assert.ok(b200.info().devices >= 1)
})

test('NAND / XOR / MUX batches decrypt to the truth table', () => {
  const key = b200.uploadCloudKey(fx('cloud.key'))
  const { n } = b200.keyParams(key)
  const bits = JSON.parse(fx('bits.json'))
  const s = JSON.parse(fx('lwe_key.json'))
  const [a, b, c] = ['a.ct', 'b.ct', 'c.ct'].map((f) => b200.readSamples(fx(f), n))
  assert.deepStrictEqual(decrypt(b200.gateBatch(key, OP.NAND, a, b), s, n), bits.a.map((x, i) => 1 - (x & bits.b[i])))
  assert.deepStrictEqual(decrypt(b200.gateBatch(key, OP.XOR, a, b), s, n), bits.a.map((x, i) => x ^ bits.b[i]))
  assert.deepStrictEqual(decrypt(b200.gateBatch(key, OP.MUX, a, b, c), s, n), bits.a.map((x, i) => (x ? bits.b[i] : bits.c[i])))
  b200.freeKey(key)
})

// ---- eoc's string entry points, case for case with tests/tfhe.test.js:142-186 (add 15 + 27, subtract) ----
test('addCiphertexts / subtractCiphertexts return the reference\'s strings', () => {
  const e = JSON.parse(fx('eoc.json'))
  assert.strictEqual(b200.addCiphertexts(e.ct15, e.ct27, ''), e.sum)        // lweCopy + lweAddTo, byte for byte
  assert.strictEqual(b200.subtractCiphertexts(e.ct27, e.ct15, ''), e.diff)  // lweCopy + lweSubTo
  assert.strictEqual(b200.addCiphertexts('AAAA', e.ct27, ''), null)         // nullptr -> null, as the Lua side sees nil
})

test('bit-sliced integers add and subtract through the bootstrapped adder', () => {
  const e = JSON.parse(fx('eoc.json'))
  const s = JSON.parse(fx('lwe_key.json'))
  b200.setCloudKeyBase64(e.cloud_key)
  const value = (b64) => decrypt(b200.readSamples(Buffer.from(b64, 'base64'), e.n), s, e.n)
    .reduce((acc, bit, i) => acc + bit * 2 ** i, 0)
  assert.strictEqual(value(b200.addCiphertextBits(e.x_bits, e.y_bits, e.nb_bits)), e.x + e.y)
  const mod = 2 ** e.nb_bits
  assert.strictEqual(value(b200.subtractCiphertextBits(e.x_bits, e.y_bits, e.nb_bits)),
    ((e.x - e.y) % mod + mod) % mod + (e.x >= e.y ? mod : 0))
})
