// Same shape as snarkjs.groth16: `const { groth16 } = require("./bindings/napi")` in place of `snarkjs.groth16` at
// git_modules/circom-ecdsa/scripts/groupsign_cli.ts:114, so snarkjs.groth16.verify(vkey, publicSignals, proof) accepts
// the result unchanged.  Unavailable on this box (no node): see g16_napi.cc.
const native = require("./build/Release/g16b200.node");

const groth16 = {
  async prove(zkeyFileName, wtnsFileName) {
    const r = native.prove(zkeyFileName, wtnsFileName);
    return { proof: JSON.parse(r.proof), publicSignals: JSON.parse(r.publicSignals) };
  },
  async verify(vkey, publicSignals, proof) {
    return native.verify(JSON.stringify(vkey), JSON.stringify(publicSignals), JSON.stringify(proof));
  },
};

module.exports = { groth16 };
