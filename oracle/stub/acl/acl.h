// Stub for <acl/acl.h>. The reference's op_tiling/utils.h:16 includes it but the selector sources use
// nothing from it. TEST INFRASTRUCTURE ONLY (oracle/_ref build); never part of the product.
#pragma once
