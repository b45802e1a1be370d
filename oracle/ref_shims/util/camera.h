// ORACLE shim: framework/render/camera.h includes util/camera.h (DirectXMath); not needed on host.
#pragma once
