// config.cu — the configs/*.yaml schema of the reference behind the C ABI (host code only).
//
// Mirror of `Config load_config(const std::filesystem::path &)` (src/bin/train.cc:108-136): the same
// 24 flat `key: value` entries with the same inline defaults.  `action_size` appears in the shipped
// YAML files but the reference never reads it (`const size_t action_size = 4`, src/bin/train.cc:36);
// it is accepted and ignored here too.  yaml-cpp is not part of this library: the shipped files are
// flat mappings of scalars, so a line parser (comments, blank lines, quoted scalars) covers the schema.
// Error behaviour follows yaml-cpp's exception types: a missing file (YAML::BadFile) or a scalar
// that does not convert (YAML::BadConversion) are std::runtime_error -> ALEGPU_RUNTIME_ERROR.
//
// alegpu_config_validate is SURVEY D6: `mini_batch_size` only sizes the Metrics tensors in the
// reference (src/bin/train.cc:383-387) while the real minibatch is E*T / num_mini_batches
// (src/ai/ppo/train.h:139-144); a file where the two disagree would silently diverge there and is
// rejected here.
#include <algorithm>
#include <cctype>
#include <cerrno>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>

#include "common.cuh"

using namespace alegpu;

namespace {

std::string trim(const std::string &s) {
  size_t a = 0, b = s.size();
  while (a < b && std::isspace((unsigned char)s[a]))
    ++a;
  while (b > a && std::isspace((unsigned char)s[b - 1]))
    --b;
  return s.substr(a, b - a);
}

// value part of a `key: value` line: strips a trailing comment (" #...") and one level of quotes
std::string scalar_of(const std::string &raw) {
  std::string v = raw;
  bool in_s = false, in_d = false;
  for (size_t i = 0; i < v.size(); ++i) {
    if (v[i] == '\'' && !in_d)
      in_s = !in_s;
    else if (v[i] == '"' && !in_s)
      in_d = !in_d;
    else if (v[i] == '#' && !in_s && !in_d && (i == 0 || std::isspace((unsigned char)v[i - 1]))) {
      v = v.substr(0, i);
      break;
    }
  }
  v = trim(v);
  if (v.size() >= 2 && ((v.front() == '"' && v.back() == '"') || (v.front() == '\'' && v.back() == '\'')))
    v = v.substr(1, v.size() - 2);
  return v;
}

bool to_i64(const std::string &v, int64_t *out) {
  if (v.empty())
    return false;
  errno = 0;
  char *end = nullptr;
  const long long x = std::strtoll(v.c_str(), &end, 10);
  if (errno || *end)
    return false;
  *out = x;
  return true;
}
bool to_u64(const std::string &v, int64_t *out) { // the reference's size_t keys: no negatives
  return to_i64(v, out) && *out >= 0;
}
bool to_f64(const std::string &v, double *out) {
  if (v.empty())
    return false;
  errno = 0;
  char *end = nullptr;
  const double x = std::strtod(v.c_str(), &end);
  if (errno || *end)
    return false;
  *out = x;
  return true;
}
bool to_bool(const std::string &v, int *out) { // yaml-cpp's bool conversion
  std::string l = v;
  std::transform(l.begin(), l.end(), l.begin(), [](unsigned char c) { return (char)std::tolower(c); });
  if (l == "true" || l == "yes" || l == "on" || l == "y") {
    *out = 1;
    return true;
  }
  if (l == "false" || l == "no" || l == "off" || l == "n") {
    *out = 0;
    return true;
  }
  return false;
}

int bad(const std::string &key, const std::string &v) {
  return fail(ALEGPU_RUNTIME_ERROR, "config: bad conversion of `" + key + ": " + v + "`");
}

int parse_text(const std::string &text, alegpu_config *c) {
  std::istringstream in(text);
  std::string line;
  while (std::getline(in, line)) {
    const std::string t = trim(line);
    if (t.empty() || t[0] == '#' || t == "---" || t == "...")
      continue;
    const size_t colon = t.find(':');
    if (colon == std::string::npos)
      return fail(ALEGPU_RUNTIME_ERROR, "config: not a `key: value` line: " + t);
    const std::string key = trim(t.substr(0, colon)), v = scalar_of(t.substr(colon + 1));
    double d = 0.0;
    int64_t i = 0;
    int b = 0;
#define KEY_U64(name)                                                                              \
  if (key == #name) {                                                                              \
    if (!to_u64(v, &i))                                                                            \
      return bad(key, v);                                                                          \
    c->name = i;                                                                                   \
    continue;                                                                                      \
  }
#define KEY_I64(name)                                                                              \
  if (key == #name) {                                                                              \
    if (!to_i64(v, &i))                                                                            \
      return bad(key, v);                                                                          \
    c->name = i;                                                                                   \
    continue;                                                                                      \
  }
#define KEY_F32(name)                                                                              \
  if (key == #name) {                                                                              \
    if (!to_f64(v, &d))                                                                            \
      return bad(key, v);                                                                          \
    c->name = (float)d;                                                                            \
    continue;                                                                                      \
  }
#define KEY_BOOL(name)                                                                             \
  if (key == #name) {                                                                              \
    if (!to_bool(v, &b))                                                                           \
      return bad(key, v);                                                                          \
    c->name = b;                                                                                   \
    continue;                                                                                      \
  }
    KEY_U64(total_environments) KEY_U64(hidden_size) KEY_U64(horizon) KEY_U64(max_steps)
    KEY_U64(frame_stack)
    if (key == "learning_rate") {
      if (!to_f64(v, &d))
        return bad(key, v);
      c->learning_rate = d;
      continue;
    }
    KEY_F32(clip_param) KEY_F32(value_loss_coef) KEY_F32(entropy_coef)
    KEY_I64(num_epochs) KEY_I64(mini_batch_size) KEY_I64(num_mini_batches)
    KEY_F32(gae_discount) KEY_F32(gae_lambda) KEY_F32(max_gradient_norm)
    KEY_U64(num_rollouts) KEY_U64(num_workers) KEY_U64(worker_batch_size) KEY_U64(frame_skip)
    KEY_F32(max_return)
    KEY_BOOL(record_observation) KEY_BOOL(record_video) KEY_BOOL(cuda_graph) KEY_BOOL(deterministic)
    KEY_U64(action_size) // present in the shipped files, never read by the reference (train.cc:36)
    // opt-in extras of this library (absent from the reference schema; default = reference semantics)
    KEY_BOOL(normalize_advantages) KEY_BOOL(clip_value_loss) KEY_BOOL(shuffle) KEY_U64(num_gpus)
#undef KEY_U64
#undef KEY_I64
#undef KEY_F32
#undef KEY_BOOL
    // unknown keys are ignored, as node["..."] lookups ignore them in the reference
  }
  return ALEGPU_OK;
}

} // namespace

// the inline defaults of src/bin/train.cc:111-134
extern "C" int alegpu_config_defaults(alegpu_config *c) {
  ALEGPU_REQUIRE(c, ALEGPU_INVALID_ARGUMENT, "config_defaults: null argument");
  std::memset(c, 0, sizeof(*c));
  c->total_environments = 512;
  c->hidden_size = 512;
  c->horizon = 128;
  c->max_steps = 108000;
  c->frame_stack = 4;
  c->learning_rate = 2.5e-4;
  c->clip_param = 0.1f;
  c->value_loss_coef = 0.5f;
  c->entropy_coef = 0.01f;
  c->num_epochs = 1;
  c->mini_batch_size = 2048;
  c->num_mini_batches = 32;
  c->gae_discount = 0.99f;
  c->gae_lambda = 0.95f;
  c->max_gradient_norm = 0.5f;
  c->num_rollouts = 7000;
  c->num_workers = 16;
  c->worker_batch_size = 32;
  c->frame_skip = 4;
  c->max_return = -1.0f;
  c->record_observation = 0;
  c->record_video = 0;
  c->cuda_graph = 0;
  c->deterministic = 0;
  c->action_size = 4; // src/bin/train.cc:36
  c->num_gpus = 1;
  return ALEGPU_OK;
}

extern "C" int alegpu_config_parse(const char *yaml_text, alegpu_config *out) {
  ALEGPU_REQUIRE(yaml_text && out, ALEGPU_INVALID_ARGUMENT, "config_parse: null argument");
  alegpu_config c;
  alegpu_config_defaults(&c);
  if (int rc = parse_text(yaml_text, &c))
    return rc;
  c.action_size = 4; // whatever the file says (src/bin/train.cc:36)
  *out = c;
  return ALEGPU_OK;
}

extern "C" int alegpu_config_load(const char *path, alegpu_config *out) {
  ALEGPU_REQUIRE(path && out, ALEGPU_INVALID_ARGUMENT, "config_load: null argument");
  std::ifstream f(path);
  if (!f.good()) // YAML::BadFile is a std::runtime_error
    return fail(ALEGPU_RUNTIME_ERROR, std::string("config: bad file: ") + path);
  std::stringstream ss;
  ss << f.rdbuf();
  return alegpu_config_parse(ss.str().c_str(), out);
}

extern "C" int alegpu_config_validate(const alegpu_config *c) {
  ALEGPU_REQUIRE(c, ALEGPU_INVALID_ARGUMENT, "config_validate: null argument");
  ALEGPU_REQUIRE(c->total_environments >= 1 && c->horizon >= 1 && c->hidden_size >= 1,
                 ALEGPU_INVALID_ARGUMENT, "config: total_environments, horizon and hidden_size must be >= 1");
  ALEGPU_REQUIRE(c->frame_stack == 4, ALEGPU_INVALID_ARGUMENT,
                 "config: frame_stack must be 4 (the network's input planes, src/bin/train.cc:232)");
  ALEGPU_REQUIRE(c->num_epochs >= 0 && c->num_mini_batches >= 1, ALEGPU_INVALID_ARGUMENT,
                 "config: num_epochs >= 0 and num_mini_batches >= 1");
  const int64_t n = c->total_environments * c->horizon;
  ALEGPU_REQUIRE(n % c->num_mini_batches == 0, ALEGPU_RUNTIME_ERROR,
                 "Batch size must be divisible by num_mini_batches"); // train.h:140-143
  ALEGPU_REQUIRE(c->mini_batch_size * c->num_mini_batches == n, ALEGPU_INVALID_ARGUMENT,
                 "config: mini_batch_size * num_mini_batches (" +
                     std::to_string(c->mini_batch_size * c->num_mini_batches) +
                     ") != total_environments * horizon (" + std::to_string(n) + ")");
  const int64_t g = c->num_gpus < 1 ? 1 : c->num_gpus;
  ALEGPU_REQUIRE((n / c->num_mini_batches) % g == 0, ALEGPU_INVALID_ARGUMENT,
                 "config: num_gpus must divide the minibatch (every rank holds mb / num_gpus rows of it)");
  return ALEGPU_OK;
}
