// DeviceB200 -- the B200-native OVR rendering device: a new back-end behind the unchanged ovr::MainRenderer plugin API.
//
// Compiled against the reference's own headers (ovr/renderer.h, ovr/scene.h, ovr/common/dylink/ObjectFactory.h) into
// libdevice_b200.so. `create_renderer("b200")` (ovr/renderer.cpp:42-61) dlopens "libdevice_b200.so" and resolves
// `ovr_create_renderer__b200` (ovr/common/dylink/ObjectFactory.h:35-85), so apps/main_batch.cpp and apps/main_app.cpp
// drive this device with `--device b200` and zero edits to reference files.
//
// Same shape as the reference devices (ovr/devices/optix7/device.h:27-55, ovr/devices/ospray/device.h): the five
// MainRenderer virtuals forwarded to a private Impl; the Impl owns one handle of the thin C ABI (include/b200vr.h)
// underneath which the hand-written sm_100a kernels live. No OptiX, no CPU fallback.
#pragma once
#ifndef OVR_B200_DEVICE_H
#define OVR_B200_DEVICE_H

#include "ovr/renderer.h"

#include <memory>

namespace ovr::b200 {

class DeviceB200 : public MainRenderer {
public:
  ~DeviceB200() override;
  DeviceB200();
  DeviceB200(const DeviceB200& other) = delete;
  DeviceB200(DeviceB200&& other) = delete;
  DeviceB200& operator=(const DeviceB200& other) = delete;
  DeviceB200& operator=(DeviceB200&& other) = delete;

  /*! builds all device state from MainRenderer::current_scene, then behaves as if commit() ran
      (ovr/devices/optix7/device.cpp:16-20) */
  void init(int argc, const char** argv) override;

  void swap() override;
  void commit() override;
  void render() override;
  void mapframe(FrameBufferData* fb) override;

private:
  struct Impl;
  std::unique_ptr<Impl> pimpl;
};

} // namespace ovr::b200
#endif // OVR_B200_DEVICE_H
