// proto2 wire codec + schema describer for the structs in pdlp_schema.h.
// Wire format facts used: tag = (field_number << 3) | wire_type; wire types 0 varint,
// 1 fixed64, 2 length-delimited, 5 fixed32; negative int32/enum values are sign-extended
// to 64 bits; packed repeated scalars are length-delimited, and a reader must also accept
// the unpacked form.
#include "pdlp_schema.h"

#include <cstdio>

namespace pdlp_b200 {

template <> const char* OptEnum<OptimalityNorm>::TypeName() { return "OptimalityNorm"; }
template <> const char* OptEnum<RestartStrategy>::TypeName() { return "PrimalDualHybridGradientParams.RestartStrategy"; }
template <> const char* OptEnum<LinesearchRule>::TypeName() { return "PrimalDualHybridGradientParams.LinesearchRule"; }
template <> const char* OptEnum<RestartChoice>::TypeName() { return "RestartChoice"; }
template <> const char* OptEnum<PointType>::TypeName() { return "PointType"; }
template <> const char* OptEnum<TerminationReason>::TypeName() { return "TerminationReason"; }
template <> const char* OptEnum<PolishingPhaseType>::TypeName() { return "PolishingPhaseType"; }

namespace {

struct EnumValue { const char* name; int number; };
struct EnumDesc { const char* name; const char* file; std::vector<EnumValue> values; };

const std::vector<EnumDesc>& Enums() {
  static const std::vector<EnumDesc> kEnums = {
      {"OptimalityNorm", "solvers", {{"OPTIMALITY_NORM_UNSPECIFIED", 0}, {"OPTIMALITY_NORM_L_INF", 1}, {"OPTIMALITY_NORM_L2", 2}, {"OPTIMALITY_NORM_L_INF_COMPONENTWISE", 3}}},
      {"PrimalDualHybridGradientParams.RestartStrategy", "solvers",
       {{"RESTART_STRATEGY_UNSPECIFIED", 0}, {"NO_RESTARTS", 1}, {"EVERY_MAJOR_ITERATION", 2}, {"ADAPTIVE_HEURISTIC", 3}, {"ADAPTIVE_DISTANCE_BASED", 4}}},
      {"PrimalDualHybridGradientParams.LinesearchRule", "solvers",
       {{"LINESEARCH_RULE_UNSPECIFIED", 0}, {"ADAPTIVE_LINESEARCH_RULE", 1}, {"MALITSKY_POCK_LINESEARCH_RULE", 2}, {"CONSTANT_STEP_SIZE_RULE", 3}}},
      {"RestartChoice", "solve_log",
       {{"RESTART_CHOICE_UNSPECIFIED", 0}, {"RESTART_CHOICE_NO_RESTART", 1}, {"RESTART_CHOICE_WEIGHTED_AVERAGE_RESET", 2}, {"RESTART_CHOICE_RESTART_TO_AVERAGE", 3}}},
      {"PointType", "solve_log",
       {{"POINT_TYPE_UNSPECIFIED", 0}, {"POINT_TYPE_CURRENT_ITERATE", 1}, {"POINT_TYPE_ITERATE_DIFFERENCE", 2}, {"POINT_TYPE_AVERAGE_ITERATE", 3}, {"POINT_TYPE_NONE", 4},
        {"POINT_TYPE_PRESOLVER_SOLUTION", 5}, {"POINT_TYPE_FEASIBILITY_POLISHING_SOLUTION", 6}}},
      {"TerminationReason", "solve_log",
       {{"TERMINATION_REASON_UNSPECIFIED", 0}, {"TERMINATION_REASON_OPTIMAL", 1}, {"TERMINATION_REASON_PRIMAL_INFEASIBLE", 2}, {"TERMINATION_REASON_DUAL_INFEASIBLE", 3},
        {"TERMINATION_REASON_TIME_LIMIT", 4}, {"TERMINATION_REASON_ITERATION_LIMIT", 5}, {"TERMINATION_REASON_KKT_MATRIX_PASS_LIMIT", 8},
        {"TERMINATION_REASON_INTERRUPTED_BY_USER", 12}, {"TERMINATION_REASON_NUMERICAL_ERROR", 6}, {"TERMINATION_REASON_INVALID_PROBLEM", 9},
        {"TERMINATION_REASON_INVALID_INITIAL_SOLUTION", 13}, {"TERMINATION_REASON_INVALID_PARAMETER", 10}, {"TERMINATION_REASON_OTHER", 7},
        {"TERMINATION_REASON_PRIMAL_OR_DUAL_INFEASIBLE", 11}}},
      {"PolishingPhaseType", "solve_log",
       {{"POLISHING_PHASE_TYPE_UNSPECIFIED", 0}, {"POLISHING_PHASE_TYPE_PRIMAL_FEASIBILITY", 1}, {"POLISHING_PHASE_TYPE_DUAL_FEASIBILITY", 2}}},
  };
  return kEnums;
}

const char* EnumValueName(const char* enum_name, int number) {
  for (const EnumDesc& e : Enums())
    if (std::strcmp(e.name, enum_name) == 0)
      for (const EnumValue& v : e.values)
        if (v.number == number) return v.name;
  return "UNKNOWN";
}

uint64_t DoubleBits(double d) { uint64_t b; std::memcpy(&b, &d, 8); return b; }
double BitsDouble(uint64_t b) { double d; std::memcpy(&d, &b, 8); return d; }

// ---------------- writer ----------------
struct Writer {
  std::string out;
  void varint(uint64_t v) { while (v >= 0x80) { out.push_back(static_cast<char>((v & 0x7f) | 0x80)); v >>= 7; } out.push_back(static_cast<char>(v)); }
  void tag(int n, int wt) { varint((static_cast<uint64_t>(n) << 3) | wt); }
  void fixed64(uint64_t b) { for (int i = 0; i < 8; ++i) out.push_back(static_cast<char>((b >> (8 * i)) & 0xff)); }
  void bytes(int n, const std::string& s) { tag(n, 2); varint(s.size()); out += s; }
  void oneof_begin(const char*) {}
  void oneof_end() {}
  void f(int n, const char*, Opt<double>& x) { if (x.has) { tag(n, 1); fixed64(DoubleBits(x.v)); } }
  void f(int n, const char*, Opt<int32_t>& x) { if (x.has) { tag(n, 0); varint(static_cast<uint64_t>(static_cast<int64_t>(x.v))); } }
  void f(int n, const char*, Opt<int64_t>& x) { if (x.has) { tag(n, 0); varint(static_cast<uint64_t>(x.v)); } }
  void f(int n, const char*, Opt<bool>& x) { if (x.has) { tag(n, 0); varint(x.v ? 1 : 0); } }
  void f(int n, const char*, Opt<std::string>& x) { if (x.has) bytes(n, x.v); }
  template <typename E> void f(int n, const char*, OptEnum<E>& x) { if (x.has) { tag(n, 0); varint(static_cast<uint64_t>(static_cast<int64_t>(x.v))); } }
  template <typename M> void f(int n, const char*, OptMsg<M>& x) { if (x.has) { Writer w; x.v.Visit(w); bytes(n, w.out); } }
  template <typename M> void f(int n, const char*, Rep<M>& xs) { for (M& m : xs) { Writer w; m.Visit(w); bytes(n, w.out); } }
  void f(int n, const char*, Packed<int32_t>& x) { if (x.v.empty()) return; Writer w; for (int32_t e : x.v) w.varint(static_cast<uint64_t>(static_cast<int64_t>(e))); bytes(n, w.out); }
  void f(int n, const char*, Packed<double>& x) { if (x.v.empty()) return; Writer w; for (double e : x.v) w.fixed64(DoubleBits(e)); bytes(n, w.out); }
  void f(int n, const char*, RawMsg& x) { if (x.has) bytes(n, x.bytes); }
};

// ---------------- reader ----------------
struct Cursor {
  const uint8_t* p; const uint8_t* end; bool ok = true;
  bool varint(uint64_t* v) {
    uint64_t r = 0;
    for (int shift = 0; shift < 64; shift += 7) {
      if (p >= end) return ok = false;
      const uint8_t b = *p++;
      r |= static_cast<uint64_t>(b & 0x7f) << shift;
      if (!(b & 0x80)) { *v = r; return true; }
    }
    return ok = false;
  }
  bool fixed64(uint64_t* v) { if (end - p < 8) return ok = false; uint64_t r = 0; for (int i = 0; i < 8; ++i) r |= static_cast<uint64_t>(p[i]) << (8 * i); p += 8; *v = r; return true; }
  bool skip(int wt) {
    uint64_t t;
    switch (wt) {
      case 0: return varint(&t);
      case 1: return fixed64(&t);
      case 2: if (!varint(&t) || static_cast<uint64_t>(end - p) < t) return ok = false; p += t; return true;
      case 5: if (end - p < 4) return ok = false; p += 4; return true;
      default: return ok = false;
    }
  }
};

template <typename M> bool ParseMessage(const uint8_t* data, size_t len, M* out);

// Visits every declared field looking for the one whose number matches the current tag.
struct FieldSetter {
  int number, wire_type;
  uint64_t scalar = 0;            // varint or fixed64 payload
  const uint8_t* data = nullptr;  // length-delimited payload
  size_t len = 0;
  bool matched = false, ok = true;
  bool in_oneof = false;
  std::vector<bool*> oneof_members;
  bool* oneof_set = nullptr;
  void oneof_begin(const char*) { in_oneof = true; oneof_members.clear(); oneof_set = nullptr; }
  void oneof_end() {
    if (oneof_set != nullptr) for (bool* b : oneof_members) if (b != oneof_set) *b = false;
    in_oneof = false;
  }
  bool hit(int n) { if (n != number) return false; matched = true; return true; }
  void f(int n, const char*, Opt<double>& x) { if (hit(n)) { if (wire_type == 1) x.set(BitsDouble(scalar)); else ok = false; } }
  void f(int n, const char*, Opt<int32_t>& x) { if (hit(n)) { if (wire_type == 0) x.set(static_cast<int32_t>(scalar)); else ok = false; } }
  void f(int n, const char*, Opt<int64_t>& x) { if (hit(n)) { if (wire_type == 0) x.set(static_cast<int64_t>(scalar)); else ok = false; } }
  void f(int n, const char*, Opt<bool>& x) { if (hit(n)) { if (wire_type == 0) x.set(scalar != 0); else ok = false; } }
  void f(int n, const char*, Opt<std::string>& x) { if (hit(n)) { if (wire_type == 2) x.set(std::string(reinterpret_cast<const char*>(data), len)); else ok = false; } }
  template <typename E> void f(int n, const char*, OptEnum<E>& x) { if (hit(n)) { if (wire_type == 0) x.set(static_cast<int>(scalar)); else ok = false; } }
  template <typename M> void f(int n, const char*, OptMsg<M>& x) {
    if (in_oneof) oneof_members.push_back(&x.has);
    if (!hit(n)) return;
    if (wire_type != 2) { ok = false; return; }
    // Repeated occurrences of a singular sub-message merge (proto2).
    x.has = true;
    if (!ParseMessage(data, len, &x.v)) ok = false;
    if (in_oneof) oneof_set = &x.has;
  }
  template <typename M> void f(int n, const char*, Rep<M>& xs) {
    if (!hit(n)) return;
    if (wire_type != 2) { ok = false; return; }
    xs.emplace_back();
    if (!ParseMessage(data, len, &xs.back())) ok = false;
  }
  void f(int n, const char*, Packed<int32_t>& x) {
    if (!hit(n)) return;
    if (wire_type == 0) { x.v.push_back(static_cast<int32_t>(scalar)); return; }
    if (wire_type != 2) { ok = false; return; }
    Cursor c{data, data + len};
    while (c.p < c.end) { uint64_t v; if (!c.varint(&v)) { ok = false; return; } x.v.push_back(static_cast<int32_t>(v)); }
  }
  void f(int n, const char*, Packed<double>& x) {
    if (!hit(n)) return;
    if (wire_type == 1) { x.v.push_back(BitsDouble(scalar)); return; }
    if (wire_type != 2) { ok = false; return; }
    Cursor c{data, data + len};
    while (c.p < c.end) { uint64_t v; if (!c.fixed64(&v)) { ok = false; return; } x.v.push_back(BitsDouble(v)); }
  }
  void f(int n, const char*, RawMsg& x) { if (hit(n)) { if (wire_type == 2) { x.has = true; x.bytes.append(reinterpret_cast<const char*>(data), len); } else ok = false; } }
};

template <typename M> bool ParseMessage(const uint8_t* data, size_t len, M* out) {
  Cursor c{data, data + len};
  while (c.p < c.end) {
    uint64_t tag;
    if (!c.varint(&tag)) return false;
    FieldSetter fs;
    fs.number = static_cast<int>(tag >> 3);
    fs.wire_type = static_cast<int>(tag & 7);
    const uint8_t* before = c.p;
    if (fs.wire_type == 0) { if (!c.varint(&fs.scalar)) return false; }
    else if (fs.wire_type == 1) { if (!c.fixed64(&fs.scalar)) return false; }
    else if (fs.wire_type == 2) {
      uint64_t l;
      if (!c.varint(&l) || static_cast<uint64_t>(c.end - c.p) < l) return false;
      fs.data = c.p; fs.len = l; c.p += l;
    } else { c.p = before; if (!c.skip(fs.wire_type)) return false; continue; }
    out->Visit(fs);
    if (!fs.ok) return false;  // wrong wire type for a known field
  }
  return true;
}

// ---------------- describer ----------------
struct Describer {
  std::string out;
  const char* oneof = "-";
  void oneof_begin(const char* name) { oneof = name; }
  void oneof_end() { oneof = "-"; }
  static std::string Dbl(double d) {
    if (std::isinf(d)) return d > 0 ? "inf" : "-inf";
    char b[40]; std::snprintf(b, sizeof b, "%.17g", d); return b;
  }
  void line(int n, const char* name, const char* label, const char* type, const std::string& type_name, const std::string& def, bool packed) {
    out += "field " + std::to_string(n) + " " + name + " " + label + " " + type + " " + (type_name.empty() ? "-" : type_name) + " " + (def.empty() ? "-" : def) + " " +
           (packed ? "1" : "0") + " " + oneof + "\n";
  }
  void f(int n, const char* nm, Opt<double>& x) { line(n, nm, "optional", "double", "", x.v != 0.0 ? Dbl(x.v) : "", false); }
  void f(int n, const char* nm, Opt<int32_t>& x) { line(n, nm, "optional", "int32", "", x.v != 0 ? std::to_string(x.v) : "", false); }
  void f(int n, const char* nm, Opt<int64_t>& x) { line(n, nm, "optional", "int64", "", x.v != 0 ? std::to_string(x.v) : "", false); }
  void f(int n, const char* nm, Opt<bool>& x) { line(n, nm, "optional", "bool", "", x.v ? "true" : "", false); }
  void f(int n, const char* nm, Opt<std::string>&) { line(n, nm, "optional", "string", "", "", false); }
  template <typename E> void f(int n, const char* nm, OptEnum<E>& x) {
    line(n, nm, "optional", "enum", OptEnum<E>::TypeName(), x.v != 0 ? EnumValueName(OptEnum<E>::TypeName(), x.v) : "", false);
  }
  template <typename M> void f(int n, const char* nm, OptMsg<M>&) { line(n, nm, "optional", "message", M::kName, "", false); }
  template <typename M> void f(int n, const char* nm, Rep<M>&) { line(n, nm, "repeated", "message", M::kName, "", false); }
  void f(int n, const char* nm, Packed<int32_t>&) { line(n, nm, "repeated", "int32", "", "", true); }
  void f(int n, const char* nm, Packed<double>&) { line(n, nm, "repeated", "double", "", "", true); }
  void f(int n, const char* nm, RawMsg& x) { line(n, nm, "optional", "message", x.type_name, "", false); }
};

template <typename M> void DescribeMessage(const char* file, std::string* out) {
  M m;
  Describer d;
  m.Visit(d);
  *out += std::string("message ") + M::kName + " " + file + "\n" + d.out + "end\n";
}

template <typename M> std::string SerializeAny(const M& m) { Writer w; const_cast<M&>(m).Visit(w); return w.out; }

}  // namespace

const char* TerminationReasonName(int r) { return EnumValueName("TerminationReason", r); }
const char* PointTypeName(int t) { return EnumValueName("PointType", t); }

std::string Serialize(const PrimalDualHybridGradientParams& m) { return SerializeAny(m); }
std::string Serialize(const SolveLog& m) { return SerializeAny(m); }
std::string Serialize(const IterationStats& m) { return SerializeAny(m); }
bool Parse(const void* data, size_t len, PrimalDualHybridGradientParams* out) { return ParseMessage(static_cast<const uint8_t*>(data), len, out); }
bool Parse(const void* data, size_t len, SolveLog* out) { return ParseMessage(static_cast<const uint8_t*>(data), len, out); }
bool Parse(const void* data, size_t len, IterationStats* out) { return ParseMessage(static_cast<const uint8_t*>(data), len, out); }

std::string DescribeSchema() {
  std::string out;
  for (const EnumDesc& e : Enums()) {
    out += std::string("enum ") + e.name + " " + e.file + "\n";
    for (const EnumValue& v : e.values) out += std::string("value ") + v.name + " " + std::to_string(v.number) + "\n";
    out += "end\n";
  }
  // Nested messages are declared before their parents so a single pass can build them.
  DescribeMessage<SimpleOptimalityCriteria>("solvers", &out);
  DescribeMessage<DetailedOptimalityCriteria>("solvers", &out);
  DescribeMessage<TerminationCriteria>("solvers", &out);
  DescribeMessage<AdaptiveLinesearchParams>("solvers", &out);
  DescribeMessage<MalitskyPockParams>("solvers", &out);
  DescribeMessage<PresolveOptions>("solvers", &out);
  DescribeMessage<PrimalDualHybridGradientParams>("solvers", &out);
  DescribeMessage<QuadraticProgramStats>("solve_log", &out);
  DescribeMessage<ConvergenceInformation>("solve_log", &out);
  DescribeMessage<InfeasibilityInformation>("solve_log", &out);
  DescribeMessage<PointMetadata>("solve_log", &out);
  DescribeMessage<IterationStats>("solve_log", &out);
  DescribeMessage<FeasibilityPolishingDetails>("solve_log", &out);
  DescribeMessage<SolveLog>("solve_log", &out);
  return out;
}

}  // namespace pdlp_b200
