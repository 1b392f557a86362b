// stream_server.hpp — frames of the B200 splat core streamed to `remote_render_ui` over packetcomms
// (BASELINE north_star: "frames still streamed to remote_render_ui over packetcomms").
//
// B200StreamServer is the reference's InterfaceServer (include/remote_ui/InterfaceServer.hpp:86-406)
// without Poplar's logger, OpenCV and FFmpeg: the same 16 packet types in the same order (:24-43, the
// order defines the wire ids), the same `ready` handshake (external/packetcomms/src/Sync.cpp:13-29),
// the same State fields and the same subscriptions writing them (:110-203, fov arriving in degrees),
// `tile_histogram` as the cereal binary of the counts (:330-332), `fov` echoed at start (:291-296).
// It is built on the reference's OWN transport — external/packetcomms (PacketMuxer / PacketDemuxer /
// TcpSocket + cereal), compiled unchanged from the reference tree by apps/Makefile.
//
// One difference, stated: `render_preview`.  The reference pushes each BGR8 frame through videolib /
// FFmpeg (MPEG-4 'FMP4', 30 fps; InterfaceServer.hpp:100-108,282,322-328) and ships the encoder's
// packets.  FFmpeg is absent from this image and B200 has no NVENC, so the payload here is the RAW
// frame: cereal binary of PreviewFrame {width, height, format = 0 (BGR8), bytes}.  PreviewSink is the
// seam: an encoder-backed sink with the same push(bgr8) call restores the reference's payload.
#pragma once

#include <PacketComms.h>
#include <PacketSerialisation.h>
#include <network/TcpSocket.h>

#include <cereal/types/string.hpp>
#include <cereal/types/vector.hpp>

#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <memory>
#include <string>
#include <thread>
#include <vector>

namespace b200ui {

// InterfaceServer.hpp:24-43 — same names, same order
inline const std::vector<std::string> &packetTypes() {
  static const std::vector<std::string> types{
      "stop", "detach", "env_rotation", "env_rotation_2", "exposure", "gamma", "X", "Y", "Z", "lambda1", "lambda2",
      "fov", "render_preview", "ready", "tile_histogram", "device"};
  return types;
}

// payload of `render_preview` while no video encoder is available (see the header comment)
struct PreviewFrame {
  std::int32_t width = 0, height = 0;
  std::uint32_t format = 0;            // 0: BGR8 rows, top to bottom (the cv::Mat getFrameBuffer fills)
  std::vector<std::uint8_t> data;
};
template <typename T> void serialize(T &ar, PreviewFrame &f) { ar(f.width, f.height, f.format, f.data); }

class B200StreamServer {
 public:
  // InterfaceServer::State, :230-244 (same fields, same defaults)
  struct State {
    float envRotationDegrees = 0.f;
    float envRotationDegrees2 = 0.f;
    float exposure = 0.f;
    float gamma = 2.2f;
    float X = 640.f;
    float Y = 360.f;
    float Z = 1.f;
    float lambda1 = 1.f;
    float lambda2 = 1.f;
    float fov = 90.f;
    std::string device = "cpu";
    bool stop = false;
    bool detach = false;
  };

  explicit B200StreamServer(int portNumber) : port(portNumber) {}
  virtual ~B200StreamServer() { stop(); }

  /// Launches the UI thread and blocks until a client has connected and the `ready` handshake is done.
  void start() {
    stopServer = false; serverReady = false; stateUpdated = false; failed = false;
    thread.reset(new std::thread(&B200StreamServer::communicate, this));
    while (!serverReady && !failed) std::this_thread::sleep_for(std::chrono::milliseconds(5));
  }
  void stop() {
    stopServer = true;
    if (thread) { thread->join(); thread.reset(); }
  }
  bool connected() const { return serverReady && !failed && sender && sender->ok(); }

  State consumeState() { State tmp = state; stateUpdated = false; return tmp; }   // :247-251
  const State &getState() const { return state; }
  bool stateChanged() const { return stateUpdated; }

  void initialiseVideoStream(std::size_t w, std::size_t h) { width = (int)w; height = (int)h; }   // :280-286
  void updateFov(float fovRadians) {                                                               // :291-296
    if (sender) { state.fov = fovRadians; serialise(*sender, "fov", fovRadians); }
  }
  /// sendPreviewImage (:322-328): one BGR8 frame, rows of `pitch` bytes
  void sendPreviewImage(const std::uint8_t *bgr8, int w, int h, std::size_t pitch) {
    if (!sender) return;
    PreviewFrame f;
    f.width = w; f.height = h; f.format = 0;
    f.data.resize((std::size_t)w * h * 3);
    for (int y = 0; y < h; ++y) std::copy(bgr8 + y * pitch, bgr8 + y * pitch + (std::size_t)w * 3, f.data.begin() + (std::size_t)y * w * 3);
    serialise(*sender, "render_preview", f);
  }
  void sendHistogram(const std::vector<std::uint32_t> &data) {   // :330-332
    if (sender) serialise(*sender, "tile_histogram", data);
  }

 private:
  void communicate() {   // :88-213
    std::fprintf(stderr, "[b200ui] user interface server listening on port %d\n", port);
    if (!serverSocket.Bind(port) || !serverSocket.Listen(0)) { failed = true; return; }
    while (!stopServer && !serverSocket.readyForReading(50)) {}
    if (stopServer) return;
    connection = serverSocket.Accept();
    if (!connection) { failed = true; return; }
    connection->setBlocking(false);
    PacketDemuxer receiver(*connection, packetTypes());
    sender.reset(new PacketMuxer(*connection, packetTypes()));
    auto f32 = [&](const char *name, float State::*field) {
      return receiver.subscribe(name, [this, field](const ComPacket::ConstSharedPacket &packet) {
        deserialise(packet, state.*field);
        stateUpdated = true;
      });
    };
    auto s1 = f32("env_rotation", &State::envRotationDegrees);
    auto s2 = f32("env_rotation_2", &State::envRotationDegrees2);
    auto s3 = f32("exposure", &State::exposure);
    auto s4 = f32("gamma", &State::gamma);
    auto s5 = f32("X", &State::X);
    auto s6 = f32("Y", &State::Y);
    auto s7 = f32("Z", &State::Z);
    auto s8 = f32("lambda1", &State::lambda1);
    auto s9 = f32("lambda2", &State::lambda2);
    auto s10 = receiver.subscribe("fov", [this](const ComPacket::ConstSharedPacket &packet) {
      deserialise(packet, state.fov);
      state.fov = state.fov * (M_PI / 180.f);   // the client sends degrees (:189-196)
      stateUpdated = true;
    });
    auto s11 = receiver.subscribe("detach", [this](const ComPacket::ConstSharedPacket &packet) {
      deserialise(packet, state.detach); stateUpdated = true;
    });
    auto s12 = receiver.subscribe("stop", [this](const ComPacket::ConstSharedPacket &packet) {
      deserialise(packet, state.stop); stateUpdated = true;
    });
    auto s13 = receiver.subscribe("device", [this](const ComPacket::ConstSharedPacket &packet) {
      deserialise(packet, state.device); stateUpdated = true;
    });
    syncWithClient(*sender, receiver, "ready");
    serverReady = true;
    while (!stopServer && receiver.ok()) std::this_thread::sleep_for(std::chrono::milliseconds(5));
    if (!receiver.ok()) state.stop = true;   // the client went away: the render loop ends as on `stop`
    sender.reset();
  }

  int port;
  int width = 0, height = 0;
  TcpSocket serverSocket;
  std::unique_ptr<std::thread> thread;
  std::atomic<bool> stopServer{false}, serverReady{false}, stateUpdated{false}, failed{false};
  std::unique_ptr<TcpSocket> connection;
  std::unique_ptr<PacketMuxer> sender;
  State state;
};

}  // namespace b200ui
