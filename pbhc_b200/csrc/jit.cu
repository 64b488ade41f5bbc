// jit.cu - config-specialised kernels: a kernel source of this library (embedded at build time, jit_sources.inc) is
// compiled by NVRTC for sm_100a together with a generated header of launch constants, so that shapes, shared-memory
// offsets and config flags of one env config are compile-time constants.  One compile per (source, constants) pair:
// cubins are cached on disk (<library dir>/jit_cache, or $PBHC_JIT_CACHE) and in the process.  NVRTC is loaded with
// dlopen on first use, so the library itself has no link-time dependency on it; when NVRTC (or the cache) cannot be
// had the caller falls back to the generic kernel built with the library - same source, same results.
//   PBHC_JIT=0 disables the specialised kernels, PBHC_JIT_VERBOSE=1 logs compiles, PBHC_NVRTC=<path> names the library.
#include <dlfcn.h>
#include <nvrtc.h>
#include <sys/stat.h>
#include <unistd.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include "common.cuh"

namespace pbhc {

#include "jit_sources.inc"  // kJitSources[] = {name, text} of csrc/*.cu|*.cuh and include/pbhc_b200.h as they were at build time

namespace {

struct Nvrtc {
  void* h = nullptr;
  decltype(&nvrtcCreateProgram) create = nullptr;
  decltype(&nvrtcCompileProgram) compile = nullptr;
  decltype(&nvrtcGetCUBINSize) cubin_size = nullptr;
  decltype(&nvrtcGetCUBIN) cubin = nullptr;
  decltype(&nvrtcGetProgramLogSize) log_size = nullptr;
  decltype(&nvrtcGetProgramLog) log = nullptr;
  decltype(&nvrtcDestroyProgram) destroy = nullptr;
  decltype(&nvrtcGetErrorString) err = nullptr;
  bool ok = false;
};

Nvrtc& nvrtc() {
  static Nvrtc n;
  static bool tried = false;
  if (tried) return n;
  tried = true;
  const char* env = getenv("PBHC_NVRTC");
  const char* cands[] = {env, "libnvrtc.so.12", "/usr/local/cuda/lib64/libnvrtc.so.12", "libnvrtc.so", "/usr/local/cuda/lib64/libnvrtc.so"};
  for (const char* c : cands) {
    if (!c || !*c) continue;
    n.h = dlopen(c, RTLD_NOW | RTLD_LOCAL);
    if (n.h) break;
  }
  if (!n.h) return n;
#define LOAD(field, sym) n.field = reinterpret_cast<decltype(n.field)>(dlsym(n.h, sym))
  LOAD(create, "nvrtcCreateProgram"); LOAD(compile, "nvrtcCompileProgram"); LOAD(cubin_size, "nvrtcGetCUBINSize");
  LOAD(cubin, "nvrtcGetCUBIN"); LOAD(log_size, "nvrtcGetProgramLogSize"); LOAD(log, "nvrtcGetProgramLog");
  LOAD(destroy, "nvrtcDestroyProgram"); LOAD(err, "nvrtcGetErrorString");
#undef LOAD
  n.ok = n.create && n.compile && n.cubin_size && n.cubin && n.log_size && n.log && n.destroy && n.err;
  return n;
}

uint64_t fnv(uint64_t h, const void* data, size_t n) {
  const unsigned char* p = static_cast<const unsigned char*>(data);
  for (size_t i = 0; i < n; ++i) { h ^= p[i]; h *= 0x100000001b3ULL; }
  return h;
}

std::string cache_dir() {
  const char* env = getenv("PBHC_JIT_CACHE");
  if (env && *env) return env;
  Dl_info info;
  if (dladdr(reinterpret_cast<const void*>(&cache_dir), &info) && info.dli_fname) {
    std::string p = info.dli_fname;
    const size_t s = p.rfind('/');
    return (s == std::string::npos ? std::string(".") : p.substr(0, s)) + "/jit_cache";
  }
  return "/tmp/pbhc_jit_cache";
}

bool read_file(const std::string& path, std::vector<char>& out) {
  FILE* f = fopen(path.c_str(), "rb");
  if (!f) return false;
  fseek(f, 0, SEEK_END);
  const long n = ftell(f);
  fseek(f, 0, SEEK_SET);
  out.resize(n > 0 ? (size_t)n : 0);
  const bool ok = n > 0 && fread(out.data(), 1, (size_t)n, f) == (size_t)n;
  fclose(f);
  return ok;
}

void write_file(const std::string& dir, const std::string& path, const std::vector<char>& data) {
  mkdir(dir.c_str(), 0777);
  const std::string tmp = path + ".tmp" + std::to_string((long)getpid());
  FILE* f = fopen(tmp.c_str(), "wb");
  if (!f) return;
  const bool ok = fwrite(data.data(), 1, data.size(), f) == data.size();
  fclose(f);
  if (ok) rename(tmp.c_str(), path.c_str()); else unlink(tmp.c_str());
}

struct Entry { int rc = PBHC_E_CUDA; const void* kernel = nullptr; bool loaded = false; std::vector<char> cubin; };
std::mutex g_mu;
std::unordered_map<uint64_t, Entry> g_entries;

}  // namespace

// the kernel `pbhc_<what>_spec` of source `<what>.cu` compiled with -DPBHC_SPEC and `spec_header` as "track_spec.h".
// kernel == nullptr: compile into the on-disk cache only (no CUDA context needed).  Returns 0 or PBHC_E_*.
int jit_get(const char* what, const std::string& spec_header, bool noise, const void** kernel) {
  static const bool off = [] { const char* e = getenv("PBHC_JIT"); return e && e[0] == '0'; }();
  static const bool verbose = [] { const char* e = getenv("PBHC_JIT_VERBOSE"); return e && e[0] != '0'; }();
  if (off) return PBHC_E_CUDA;
  static const bool prof = [] { const char* e = getenv("PBHC_TILE_PROF"); return e && e[0] == '1'; }();  // development: clock stamps in the tile kernel
  // csrc/track.cu is the reference's unfused fp32 operation sequence (built with -fmad=false by build.py): the same here;
  // csrc/track_tile.cu spells its bit-exact chains with __fmul_rn & co. and lets the rest contract
  const bool nofma = strcmp(what, "track") == 0;
  const std::string opts_tag = std::string("sm_100a;c++17;lineinfo;default-device;noise=") + (noise ? "1" : "0") + ";" + what + (prof ? ";prof" : "") + (nofma ? ";fmad=false" : "");
  uint64_t key = 0xcbf29ce484222325ULL;
  key = fnv(key, opts_tag.data(), opts_tag.size());
  key = fnv(key, spec_header.data(), spec_header.size());
  for (const auto& s : kJitSources) key = fnv(key, s.text, strlen(s.text));
  std::lock_guard<std::mutex> lock(g_mu);
  Entry& e = g_entries[key];
  if (e.rc == PBHC_OK && (e.loaded || !kernel)) { if (kernel) *kernel = e.kernel; return PBHC_OK; }
  if (e.rc != PBHC_OK && e.cubin.empty() && e.loaded) return e.rc;  // (a failure is remembered: loaded marks "tried")
  if (e.cubin.empty()) {
    char name[64];
    snprintf(name, sizeof(name), "/%s_%016llx.cubin", what, (unsigned long long)key);
    const std::string dir = cache_dir(), path = dir + name;
    if (!read_file(path, e.cubin)) {
      e.cubin.clear();
      Nvrtc& n = nvrtc();
      if (!n.ok) { e.loaded = true; set_error("jit: NVRTC is not available (%s)", n.h ? "missing symbols" : "dlopen failed"); return e.rc = PBHC_E_CUDA; }
      const auto t0 = std::chrono::steady_clock::now();
      std::string main_src;
      std::vector<const char*> hdr_text, hdr_name;
      const std::string main_name = std::string(what) + ".cu";
      for (const auto& s : kJitSources) {
        if (main_name == s.name) main_src = s.text;
        else { hdr_text.push_back(s.text); hdr_name.push_back(s.name); }
      }
      hdr_text.push_back(spec_header.c_str()); hdr_name.push_back("track_spec.h");
      if (main_src.empty()) { e.loaded = true; set_error("jit: source %s is not embedded", main_name.c_str()); return e.rc = PBHC_E_BADARG; }
      nvrtcProgram prog;
      if (n.create(&prog, main_src.c_str(), main_name.c_str(), (int)hdr_text.size(), hdr_text.data(), hdr_name.data()) != NVRTC_SUCCESS) {
        e.loaded = true; set_error("jit: nvrtcCreateProgram failed"); return e.rc = PBHC_E_CUDA;
      }
      const std::string dnoise = std::string("-DPBHC_NOISE=") + (noise ? "1" : "0");
      std::vector<const char*> opts = {"--gpu-architecture=sm_100a", "-std=c++17", "-lineinfo", "-default-device", "-DPBHC_SPEC=1", dnoise.c_str()};  // (-default-device: the C ABI declarations of the header are host functions)
      if (prof) opts.push_back("-DPBHC_TILE_PROF=1");
      if (nofma) opts.push_back("--fmad=false");
      const nvrtcResult r = n.compile(prog, (int)opts.size(), opts.data());
      size_t ls = 0;
      n.log_size(prog, &ls);
      std::string log(ls, '\0');
      if (ls > 1) n.log(prog, &log[0]);
      if (r != NVRTC_SUCCESS) {
        set_error("jit: NVRTC failed on %s: %s\n%.300s", main_name.c_str(), n.err(r), log.c_str());
        if (verbose) fprintf(stderr, "[pbhc jit] compile of %s FAILED: %s\n%s\n", main_name.c_str(), n.err(r), log.c_str());
        n.destroy(&prog);
        e.loaded = true;
        return e.rc = PBHC_E_CUDA;
      }
      size_t cs = 0;
      n.cubin_size(prog, &cs);
      e.cubin.resize(cs);
      n.cubin(prog, e.cubin.data());
      n.destroy(&prog);
      write_file(dir, path, e.cubin);
      if (verbose) {
        const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        fprintf(stderr, "[pbhc jit] compiled %s (%zu B cubin, %.0f ms) -> %s\n", main_name.c_str(), cs, ms, path.c_str());
      }
    } else if (verbose) fprintf(stderr, "[pbhc jit] %s from the cache: %s\n", what, path.c_str());
  }
  e.rc = PBHC_OK;
  if (!kernel) return PBHC_OK;
  cudaLibrary_t lib;
  cudaKernel_t fn;
  const std::string sym = std::string("pbhc_") + what + "_spec";
  if (cudaLibraryLoadData(&lib, e.cubin.data(), nullptr, nullptr, 0, nullptr, nullptr, 0) != cudaSuccess ||
      cudaLibraryGetKernel(&fn, lib, sym.c_str()) != cudaSuccess) {
    set_error("jit: loading the specialised %s kernel failed: %s", what, cudaGetErrorString(cudaGetLastError()));
    e.loaded = true; e.cubin.clear();
    return e.rc = PBHC_E_CUDA;
  }
  e.kernel = reinterpret_cast<const void*>(fn);
  e.loaded = true;
  *kernel = e.kernel;
  return PBHC_OK;
}

}  // namespace pbhc
