// renderloop -- the interactive render loop of the reference's GUI (apps/main_app.cpp:233-263, MainWindow::render_background)
// without a window: commit -> mapframe -> to_cpu -> swap -> render, once per iteration, against any ovr::MainRenderer back-end
// found by create_renderer (ovr/renderer.cpp:42-61). The GUI itself needs GLFW / OpenGL (unavailable on this box); what it asks
// of a device is exactly this call sequence, in which the frame handed out by mapframe is the one rendered in the PREVIOUS
// iteration and the device renders the next one into the other half of its double buffer (device_impl.cpp:102-111, 271-281).
//
// Built by device/build.py against the reference's unchanged headers and host objects (build/ovr/libovr_host.a); nothing of the
// reference is copied. Every iteration moves the camera (a slow orbit about the up axis), so every frame is a different image.
//
//   renderloop --scene s.json --device b200 --fbsize W H --frames N --out prefix [--no-swap]
//
// writes prefix<k>.rgba (W * H * 4 floats, the frame mapped in iteration k = the frame rendered in iteration k - 1; iteration 0
// maps the cleared buffer) and prints one line per iteration with the camera it committed ("%.9g": exact floats) and the host time
// of each phase.
#include "ovr/renderer.h"
#include "ovr/scene.h"
#include "ovr/serializer/serializer.h"

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

using namespace ovr;

static double
now_ms()
{
  return std::chrono::duration<double, std::milli>(std::chrono::high_resolution_clock::now().time_since_epoch()).count();
}

extern "C" int
main(int ac, char** av)
{
  std::string scene_file, device = "b200", out = "frame";
  int W = 320, H = 200, frames = 4;
  bool do_swap = true;
  for (int i = 1; i < ac; i++) {
    const std::string a = av[i];
    if (a == "--scene" && i + 1 < ac) scene_file = av[++i];
    else if (a == "--device" && i + 1 < ac) device = av[++i];
    else if (a == "--out" && i + 1 < ac) out = av[++i];
    else if (a == "--frames" && i + 1 < ac) frames = std::atoi(av[++i]);
    else if (a == "--fbsize" && i + 2 < ac) W = std::atoi(av[++i]), H = std::atoi(av[++i]);
    else if (a == "--no-swap") do_swap = false;
  }
  if (scene_file.empty()) {
    std::fprintf(stderr, "usage: renderloop --scene s.json [--device b200] [--fbsize W H] [--frames N] [--out prefix]\n");
    return 2;
  }
  Scene sc = ovr::scene::create_json_scene(scene_file); // as apps/main_batch.cpp:245
  Camera camera = sc.camera;
  auto ren = create_renderer(device);
  ren->set_fbsize(vec2i(W, H));
  ren->init(ac, (const char**)av, sc, camera);

  MainRenderer::FrameBufferData fb;
  const vec3f at = camera.at, up = camera.up;
  const vec3f d0 = camera.from - camera.at;
  for (int k = 0; k <= frames; k++) {
    // orbit about the up axis (y): 7 degrees per iteration
    const float ang = (float)k * 7.f * (float)M_PI / 180.f;
    const float c = std::cos(ang), s = std::sin(ang);
    const vec3f d(c * d0.x + s * d0.z, d0.y, -s * d0.x + c * d0.z);
    Camera cam = camera;
    cam.from = at + d;
    const double t0 = now_ms();
    ren->set_camera(cam);
    ren->commit();
    const double t1 = now_ms();
    ren->mapframe(&fb);
    const vec4f* rgba = (const vec4f*)fb.rgba->to_cpu()->data(); // frame of the previous iteration, copied to the host
    const double t2 = now_ms();
    {
      char name[1024];
      std::snprintf(name, sizeof(name), "%s%03d.rgba", out.c_str(), k);
      std::FILE* f = std::fopen(name, "wb");
      if (!f) {
        std::fprintf(stderr, "cannot write %s\n", name);
        return 1;
      }
      std::fwrite(rgba, sizeof(vec4f), (size_t)W * H, f);
      std::fclose(f);
    }
    const double t3 = now_ms();
    if (do_swap) ren->swap();
    ren->render(); // frame k, into the other buffer
    const double t4 = now_ms();
    std::printf("iteration %d camera %.9g %.9g %.9g at %.9g %.9g %.9g up %.9g %.9g %.9g fovy %.9g | ms commit %.3f map+to_cpu %.3f swap+render %.3f\n", k,
                cam.from.x, cam.from.y, cam.from.z, at.x, at.y, at.z, up.x, up.y, up.z, cam.perspective.fovy, t1 - t0, t2 - t1, t4 - t3);
  }
  return 0;
}
