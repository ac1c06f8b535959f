// N-API addon exposing snarkjs' JS API over libg16b200 -- the host binding `north_star` asks for.
//
//   const { proof, publicSignals } = await snarkjs.groth16.prove(zkeyFileName, wtnsFileName)
//       (reference call site: git_modules/circom-ecdsa/scripts/groupsign_cli.ts:114; verified right after at :123)
//
// UNAVAILABLE ON THIS BOX: the build image has no `node`, `npm` or node-addon-api headers, so this file is committed
// as source for the first machine that has them (`npm install node-addon-api && npx node-gyp rebuild` in this directory;
// binding.gyp links ../../zk-proof-of-assets_b200/libg16b200.so).  It uses only the C ABI of include/g16_b200.h, the same
// calls the Python mirror and the `prover` CLI make, which ARE built and tested here.
#include <napi.h>

#include <cstdio>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "g16_b200.h"

namespace {

std::mutex g_mu;
std::map<std::string, g16_zkey*> g_keys;   // resident keys, one per path (the workflow reuses a zkey across batches)

g16_zkey* cached_open(const std::string& path, const g16_opts* o) {
  std::lock_guard<std::mutex> lk(g_mu);
  auto it = g_keys.find(path);
  if (it != g_keys.end()) return it->second;
  g16_zkey* zk = nullptr;
  if (g16_zkey_open(path.c_str(), o, &zk) != G16_OK) return nullptr;
  g_keys[path] = zk;
  return zk;
}

bool read_file(const std::string& path, std::vector<uint8_t>& out) {
  FILE* f = std::fopen(path.c_str(), "rb");
  if (!f) return false;
  std::fseek(f, 0, SEEK_END);
  long n = std::ftell(f);
  std::fseek(f, 0, SEEK_SET);
  out.resize(n > 0 ? size_t(n) : 0);
  bool ok = n >= 0 && std::fread(out.data(), 1, out.size(), f) == out.size();
  std::fclose(f);
  return ok;
}

// prove(zkeyFileName, wtnsFileName) -> { proof: <JSON text>, publicSignals: <JSON text> }   (parsed by index.js)
Napi::Value Prove(const Napi::CallbackInfo& info) {
  Napi::Env env = info.Env();
  if (info.Length() < 2 || !info[0].IsString() || !info[1].IsString()) {
    Napi::TypeError::New(env, "prove(zkeyFileName, wtnsFileName)").ThrowAsJavaScriptException();
    return env.Null();
  }
  std::string zkey = info[0].As<Napi::String>(), wtns = info[1].As<Napi::String>();
  g16_opts o;
  g16_opts_default(&o);
  o.style = G16_STYLE_SNARKJS;
  g16_zkey* zk = cached_open(zkey, &o);
  if (!zk) {
    Napi::Error::New(env, g16_last_error()).ThrowAsJavaScriptException();
    return env.Null();
  }
  std::vector<uint8_t> w;
  if (!read_file(wtns, w)) {
    Napi::Error::New(env, "cannot read " + wtns).ThrowAsJavaScriptException();
    return env.Null();
  }
  std::vector<char> proof(4096), pub(1 << 16);
  size_t pl = proof.size(), ul = pub.size();
  int rc = g16_prove(zk, w.data(), w.size(), proof.data(), &pl, pub.data(), &ul, &o);
  if (rc == G16_ERR_SHORT_BUFFER) {   // sizes were written back
    proof.resize(pl + 1);
    pub.resize(ul + 1);
    pl = proof.size();
    ul = pub.size();
    rc = g16_prove(zk, w.data(), w.size(), proof.data(), &pl, pub.data(), &ul, &o);
  }
  if (rc != G16_OK) {
    Napi::Error::New(env, g16_last_error()).ThrowAsJavaScriptException();
    return env.Null();
  }
  Napi::Object out = Napi::Object::New(env);
  out.Set("proof", std::string(proof.data(), pl));
  out.Set("publicSignals", std::string(pub.data(), ul));
  return out;
}

// verify(vkeyJsonText, publicSignalsJsonText, proofJsonText) -> boolean   (snarkjs.groth16.verify, scripts/g16_verify.sh:212-216)
Napi::Value Verify(const Napi::CallbackInfo& info) {
  Napi::Env env = info.Env();
  if (info.Length() < 3) {
    Napi::TypeError::New(env, "verify(vkey, publicSignals, proof) takes JSON texts").ThrowAsJavaScriptException();
    return env.Null();
  }
  std::string v = info[0].As<Napi::String>(), u = info[1].As<Napi::String>(), p = info[2].As<Napi::String>();
  int ok = 0;
  if (g16_verify(v.data(), v.size(), u.data(), u.size(), p.data(), p.size(), &ok) != G16_OK) {
    Napi::Error::New(env, g16_last_error()).ThrowAsJavaScriptException();
    return env.Null();
  }
  return Napi::Boolean::New(env, ok != 0);
}

Napi::Object Init(Napi::Env env, Napi::Object exports) {
  exports.Set("prove", Napi::Function::New(env, Prove));
  exports.Set("verify", Napi::Function::New(env, Verify));
  return exports;
}

}  // namespace

NODE_API_MODULE(g16b200, Init)
