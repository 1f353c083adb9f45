// Stub for CANN's "tiling/platform/platform_ascendc.h": just enough for the reference's PlatformInfo
// (aclnn_catlass_dynamic_matmul/op_host/op_tiling/platform_info.h:16-41) to construct without CANN.
// Values: Ascend 910B defaults from the same header (coreNum 24, UB 192K, L1 512K, L0A/B 64K, L0C 128K);
// the core count can be overridden with REF_CORE_NUM (the Python mirror defaults to 20,
// get_best_config/tiling_calculator.py:22-30).
// TEST INFRASTRUCTURE ONLY (oracle/_ref build); never part of the product.
#pragma once
#include <cstdint>
#include <cstdlib>

namespace platform_ascendc {
enum class CoreMemType { UB, L1, L0_A, L0_B, L0_C };

class PlatformAscendCManager {
public:
    static PlatformAscendCManager* GetInstance() {
        static PlatformAscendCManager inst;
        return &inst;
    }
    uint32_t GetCoreNumAic() const {
        const char* env = std::getenv("REF_CORE_NUM");
        return env ? static_cast<uint32_t>(std::atoi(env)) : 24u;
    }
    void GetCoreMemSize(CoreMemType type, uint64_t& size) const {
        switch (type) {
            case CoreMemType::UB: size = 192 * 1024; break;
            case CoreMemType::L1: size = 512 * 1024; break;
            case CoreMemType::L0_A: size = 64 * 1024; break;
            case CoreMemType::L0_B: size = 64 * 1024; break;
            case CoreMemType::L0_C: size = 128 * 1024; break;
        }
    }
};
}  // namespace platform_ascendc
