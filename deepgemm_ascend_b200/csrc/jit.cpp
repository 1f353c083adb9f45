// Run-time specialisation cache for the persistent GEMM kernel.
//
// Role of the reference framework's JIT layer:
//   * code generation  - generate_code.hpp:123-369 + gemm.hpp:44-61 (a constant block formatted into a fixed text)
//   * Compiler::build  - compiler.hpp:26-52: signature "b{}m{}n{}k{}_type{}", directory
//                        $DGA_ROOT_DIR/deep_gemm_ascend/cache/kernel_<signature>/, source written next to the binary,
//                        external compiler run through the shell (compiler.hpp:77-95), runtime created from the dir
//   * KernelRuntimeCache - cache.hpp:11-33: in-process map keyed by directory, then an on-disk validity check
//                        (source and binary both present, kernel_runtime.hpp:30-33), else compile
//   * KernelRuntime    - kernel_runtime.hpp:13-21: load the binary, look the entry up by name
// B200 differences: the compiler is nvcc (-cubin for sm_100a, about a second per kernel instead of a cmake
// configure + build), the binary is loaded through the CUDA runtime's library API, the on-disk check also compares
// the stored source with what would be generated now (a stale directory from another version is rebuilt, not
// trusted), the binary appears under its final name only by rename (two processes may race for one directory), and
// the maps are mutex-guarded.
#include "jit.h"

#include <dlfcn.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <map>
#include <mutex>
#include <sstream>

namespace dga {
namespace jit {

namespace {

constexpr const char* kCodeName = "kernel.cu";
constexpr const char* kBinName = "kernel.cubin";
constexpr const char* kEntryName = "dga_jit_gemm";
constexpr int kAbiVersion = 2;   // bump when jit_entry.cuh's contract with the host changes
constexpr const char* kCompileFlags =
    " -gencode arch=compute_100a,code=sm_100a -std=c++17 -O3 -lineinfo --expt-relaxed-constexpr -cubin";
// headers the generated unit pulls in: their CONTENT is part of a cached kernel's identity
constexpr const char* kIncluded[] = {"jit_entry.cuh", "gemm_sm100.cuh", "ptx.cuh"};

std::mutex g_mu;
std::map<std::string, cudaKernel_t> g_loaded;   // kernel_dir -> kernel (libraries stay loaded for the process)
Stats g_stats;

std::string env_or(const char* name, const std::string& fallback) {
    const char* v = std::getenv(name);
    return (v != nullptr && *v != '\0') ? std::string(v) : fallback;
}

bool exists(const std::string& path) {
    struct stat st;
    return ::stat(path.c_str(), &st) == 0;
}

// Creates the missing directories of `path` readable by the owner only.
bool make_dirs(const std::string& path) {
    std::string cur;
    for (size_t i = 0; i <= path.size(); ++i) {
        if (i == path.size() || path[i] == '/') {
            if (!cur.empty() && !exists(cur) && ::mkdir(cur.c_str(), 0700) != 0 && !exists(cur)) return false;
        }
        if (i < path.size()) cur.push_back(path[i]);
    }
    return true;
}

// A cache directory is trusted only if it is a real directory owned by this user that nobody else can write to:
// binaries found there are loaded into the process (another local user must not be able to plant one).
bool trusted_dir(const std::string& path, std::string* why) {
    struct stat st;
    if (::lstat(path.c_str(), &st) != 0) { *why = "cannot stat " + path; return false; }
    if (!S_ISDIR(st.st_mode)) { *why = path + " is not a directory"; return false; }
    if (st.st_uid != ::geteuid()) { *why = path + " belongs to another user"; return false; }
    if ((st.st_mode & (S_IWGRP | S_IWOTH)) != 0) { *why = path + " is writable by other users"; return false; }
    return true;
}

std::string shell_quote(const std::string& v) {
    std::string out = "'";
    for (char c : v) {
        if (c == '\'') out += "'\\''"; else out.push_back(c);
    }
    out.push_back('\'');
    return out;
}

uint64_t fnv1a(const std::string& data, uint64_t h = 1469598103934665603ull) {
    for (unsigned char c : data) { h ^= c; h *= 1099511628211ull; }
    return h;
}

std::string read_file(const std::string& path) {
    std::ifstream f(path, std::ios::binary);
    std::stringstream ss;
    ss << f.rdbuf();
    return ss.str();
}

// Directory holding gemm_sm100.cuh / jit_entry.cuh: $DGA_JIT_INCLUDE_DIR, else ../csrc next to this library.
std::string include_dir() {
    const std::string forced = env_or("DGA_JIT_INCLUDE_DIR", "");
    if (!forced.empty()) return forced;
    Dl_info info;
    if (::dladdr(reinterpret_cast<const void*>(&include_dir), &info) != 0 && info.dli_fname != nullptr) {
        std::string lib(info.dli_fname);
        const size_t slash = lib.rfind('/');
        const std::string lib_dir = slash == std::string::npos ? "." : lib.substr(0, slash);
        return lib_dir + "/../csrc";
    }
    return "csrc";
}

std::string compiler_path() {
    const std::string forced = env_or("NVCC", "");
    if (!forced.empty()) return forced;
    if (exists("/usr/local/cuda/bin/nvcc")) return "/usr/local/cuda/bin/nvcc";
    return "nvcc";
}

const char* out_type_name(int32_t dtype) {
    switch (dtype) {
        case 0: return "float";
        case 1: return "__half";
        default: return "__nv_bfloat16";
    }
}

std::string cache_root() { return env_or("DGA_ROOT_DIR", "/tmp/dga_b200_" + std::to_string(::getuid())); }
std::string cache_dir() { return cache_root() + "/deep_gemm_ascend/cache"; }

// Runs `command` through the shell, returns its exit status and the combined output (reference:
// call_external_command, utils/system.hpp).
int run_command(const std::string& command, std::string* output) {
    FILE* pipe = ::popen((command + " 2>&1").c_str(), "r");
    if (pipe == nullptr) return -1;
    char buf[512];
    while (::fgets(buf, sizeof(buf), pipe) != nullptr) output->append(buf);
    const int status = ::pclose(pipe);
    return status;
}

// Identity of the toolchain and of the kernel text behind every generated unit: content of the included headers,
// the compiler's version banner and the flags.  Computed once per process.  A kernel directory written under any
// other digest is rebuilt (its stored source no longer equals what is generated now).
const std::string& build_digest() {
    static const std::string digest = [] {
        uint64_t h = fnv1a(kCompileFlags);
        const std::string inc = include_dir();
        for (const char* name : kIncluded) h = fnv1a(read_file(inc + "/" + name), h ^ fnv1a(name));
        std::string banner;
        run_command(shell_quote(compiler_path()) + " --version", &banner);
        h = fnv1a(banner, h);
        char buf[32];
        std::snprintf(buf, sizeof(buf), "%016llx", static_cast<unsigned long long>(h));
        return std::string(buf);
    }();
    return digest;
}

bool valid_on_disk(const std::string& dir, const std::string& code) {
    std::string why;
    return exists(dir + kCodeName) && exists(dir + kBinName) && trusted_dir(cache_dir(), &why) && trusted_dir(dir, &why) &&
           read_file(dir + kCodeName) == code;
}

bool compile(const std::string& dir, const std::string& code, std::string* err) {
    if (!make_dirs(dir)) {
        *err = "cannot create " + dir;
        return false;
    }
    if (!trusted_dir(cache_dir(), err) || !trusted_dir(dir, err)) {
        *err = "refusing the kernel cache: " + *err + " (set DGA_ROOT_DIR to a directory of your own)";
        return false;
    }
    const std::string pid = std::to_string(::getpid());
    {
        // source and binary both reach their final names by rename: a concurrent builder of the same directory
        // never reads a half-written file
        const std::string code_tmp = dir + kCodeName + ".tmp" + pid;
        std::ofstream f(code_tmp, std::ios::binary | std::ios::trunc);
        f << code;
        f.close();
        if (!f.good() || ::rename(code_tmp.c_str(), (dir + kCodeName).c_str()) != 0) {
            *err = "cannot write " + dir + kCodeName;
            ::unlink(code_tmp.c_str());
            return false;
        }
    }
    const std::string tmp = dir + kBinName + ".tmp" + pid;
    const std::string command = shell_quote(compiler_path()) + kCompileFlags + " -I" + shell_quote(include_dir()) + " -o " +
                                shell_quote(tmp) + " " + shell_quote(dir + kCodeName);
    std::string output;
    const int status = run_command(command, &output);
    if (status != 0 || !exists(tmp)) {
        if (output.size() > 600) output = output.substr(output.size() - 600);
        *err = "kernel compilation failed (" + command + "): " + output;
        ::unlink(tmp.c_str());
        return false;
    }
    if (::rename(tmp.c_str(), (dir + kBinName).c_str()) != 0) {
        *err = "cannot move the compiled kernel into " + dir;
        return false;
    }
    return true;
}

}  // namespace

std::string signature(const Request& r) {
    // variant knobs, packed: mode[0,2) cta-1[2] b_mn[3] out[4,6) in[6] block_n/16[7,12) stages[12,16)
    // raster_group[16,24) raster_along_m[24] split_k[25,30)
    const uint32_t out_code = r.out_dtype == 0 ? 2u : (r.out_dtype == 1 ? 1u : 0u);
    const uint32_t in_code = r.in_dtype == 1 ? 1u : 0u;
    const uint32_t type = (uint32_t(r.mode) & 3u) | ((r.cta - 1u) & 1u) << 2 | (r.b_mn ? 1u : 0u) << 3 | out_code << 4 |
                          in_code << 6 | ((r.block_n / 16u) & 31u) << 7 | (r.stages & 15u) << 12 |
                          (r.raster_group & 255u) << 16 | (r.raster_along_m & 1u) << 24 | (r.split_k & 31u) << 25;
    char buf[128];
    std::snprintf(buf, sizeof(buf), "b%um%un%uk%u_type%u", r.batch, r.m, r.n, r.k, type);
    return buf;
}

std::string kernel_dir(const Request& r) {
    return cache_dir() + "/kernel_" + signature(r) + "/";
}

std::string generate_code(const Request& r) {
    std::ostringstream s;
    s << "// generated by dga-b200 (jit abi " << kAbiVersion << ", build " << build_digest() << "): " << signature(r) << "\n";
    auto def = [&s](const char* name, uint64_t v) { s << "#define DGA_JIT_" << name << " " << v << "u\n"; };
    def("M", r.m);
    def("N", r.n);
    def("K", r.k);
    def("BLOCK_N", r.block_n);
    def("STAGES", r.stages);
    def("K_BLOCKS", r.k_blocks);
    def("M_BLOCKS", r.m_blocks);
    def("N_BLOCKS", r.n_blocks);
    def("RASTER_GROUP", r.raster_group);
    def("RASTER_ALONG_M", r.raster_along_m);
    def("IDESC", r.idesc);
    def("B_STAGE_BYTES", r.b_stage_bytes);
    def("K_BLOCK_ELEMS", r.k_block_elems);
    def("DP_TILES", r.dp_tiles);
    def("EPI_COLS", r.epi_cols);
    def("TILES_PER_BATCH", r.tiles_per_batch);
    def("SPLIT_K", r.split_k);
    const std::string keep = env_or("DGA_JIT_KEEP_RT", "");   // developer knob, see jit_entry.cuh
    if (!keep.empty()) s << "#define DGA_JIT_KEEP_RT " << std::atoi(keep.c_str()) << "\n";
    s << "#define DGA_JIT_CTA " << r.cta << "\n";
    s << "#define DGA_JIT_OUT_T " << out_type_name(r.out_dtype) << "\n";
    s << "#define DGA_JIT_B_MN " << (r.b_mn ? "true" : "false") << "\n";
    s << "#define DGA_JIT_MODE " << r.mode << "\n";
    s << "#include \"jit_entry.cuh\"\n";
    return s.str();
}

bool ensure_compiled(const Request& r, std::string* dir_out, std::string* err) {
    const std::string dir = kernel_dir(r);
    const std::string code = generate_code(r);
    if (dir_out != nullptr) *dir_out = dir;
    if (valid_on_disk(dir, code)) return true;
    if (!compile(dir, code, err)) return false;
    std::lock_guard<std::mutex> lock(g_mu);
    ++g_stats.compiled;
    return true;
}

bool get_kernel(const Request& r, cudaKernel_t* kernel, std::string* err) {
    const std::string dir = kernel_dir(r);
    {
        std::lock_guard<std::mutex> lock(g_mu);
        auto it = g_loaded.find(dir);
        if (it != g_loaded.end()) {
            ++g_stats.memory_hits;
            *kernel = it->second;
            return true;
        }
    }
    const uint64_t compiled_before = stats().compiled;
    if (!ensure_compiled(r, nullptr, err)) return false;
    cudaLibrary_t library = nullptr;
    cudaError_t e = cudaLibraryLoadFromFile(&library, (dir + kBinName).c_str(), nullptr, nullptr, 0, nullptr, nullptr, 0);
    if (e != cudaSuccess) {
        *err = "cannot load " + dir + kBinName + ": " + cudaGetErrorString(e);
        return false;
    }
    cudaKernel_t k = nullptr;
    e = cudaLibraryGetKernel(&k, library, kEntryName);
    if (e != cudaSuccess) {
        *err = std::string("entry ") + kEntryName + " not found in " + dir + kBinName + ": " + cudaGetErrorString(e);
        cudaLibraryUnload(library);
        return false;
    }
    std::lock_guard<std::mutex> lock(g_mu);
    auto ins = g_loaded.emplace(dir, k);
    if (!ins.second) {
        cudaLibraryUnload(library);   // another thread won the race
    } else if (g_stats.compiled == compiled_before) {
        ++g_stats.disk_hits;
    }
    *kernel = ins.first->second;
    return true;
}

Stats stats() {
    std::lock_guard<std::mutex> lock(g_mu);
    return g_stats;
}

void drop_memory_cache() {
    std::lock_guard<std::mutex> lock(g_mu);
    g_loaded.clear();   // the libraries stay loaded; only the lookups are forgotten
}

}  // namespace jit
}  // namespace dga
