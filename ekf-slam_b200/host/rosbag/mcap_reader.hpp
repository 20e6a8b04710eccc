#ifndef EKF_SLAM_B200_HOST_ROSBAG_MCAP_READER_HPP
#define EKF_SLAM_B200_HOST_ROSBAG_MCAP_READER_HPP
// rosbag::McapReader - reads the two topics the SLAM node consumes out of a rosbag2 MCAP file, without ROS:
//
//   /blue/joint_states          sensor_msgs/msg/JointState          -> Slam::jointStateCallback (slam.cpp:599-634)
//   /green/detected_obstacles   visualization_msgs/msg/MarkerArray  -> Slam::sensor_cb          (slam.cpp:318-530)
//   (nusim/fake_sensor, same type, for the known-association callback slam.cpp:180-316)
//
// SURVEY.md 8(f) row N4.  The reference records exactly these with `ros2 bag record` into
// rosbags/rosbag2_2024_03_17-18_35_57 (metadata.yaml: storage_identifier mcap, serialization_format cdr,
// 380 detection arrays, 8,419 joint states); the .mcap payload itself is not in the checkout
// (.MISSING_LARGE_BLOBS:2), so this reader is written against the published MCAP v0 container layout and the
// OMG CDR (XCDR1, little endian) encoding of the two message definitions and is PINNED ONLY by round trips through
// an independent writer (tests/_mcap_writer.py): "parity unpinned" until the blob is available.
//
// Container: magic, then records {opcode u8, length u64, body}.  The data section is walked linearly (Schema,
// Channel, Message, Chunk -> the same records inside); the summary section is not needed.  Chunk compression:
// "" always; "zstd" / "lz4" when libzstd.so.1 / liblz4.so.1 can be dlopen()ed (no headers needed), otherwise
// the open fails with a message that says so.
#include <dlfcn.h>

#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace rosbag
{

struct JointStateMsg
{
  uint64_t log_time = 0;             ///< MCAP log time, ns
  int64_t stamp_ns = 0;              ///< header.stamp
  std::vector<std::string> name;
  std::vector<double> position, velocity, effort;
};

struct MarkerMsg                      ///< the fields the callbacks read (slam.cpp:203-210, 340-347)
{
  int32_t id = 0, type = 0, action = 0;
  double x = 0.0, y = 0.0, z = 0.0;   ///< pose.position
  double scale[3] = {0.0, 0.0, 0.0};
};

struct MarkerArrayMsg
{
  uint64_t log_time = 0;
  std::vector<MarkerMsg> markers;
};

/// bounds-checked little-endian cursor over a byte range
class Cursor
{
public:
  Cursor(const uint8_t * p, size_t n)
  : p_(p), n_(n) {}
  size_t pos() const {return i_;}
  size_t left() const {return n_ - i_;}
  const uint8_t * here() const {return p_ + i_;}
  void skip(size_t k) {need(k); i_ += k;}
  template<class T> T get()
  {
    need(sizeof(T));
    T v;
    std::memcpy(&v, p_ + i_, sizeof(T));
    i_ += sizeof(T);
    return v;
  }
  std::string str32()                 ///< MCAP string: u32 byte length, no terminator
  {
    const uint32_t k = get<uint32_t>();
    need(k);
    std::string s(reinterpret_cast<const char *>(p_ + i_), k);
    i_ += k;
    return s;
  }

private:
  void need(size_t k) const
  {
    if (k > n_ - i_) {throw std::runtime_error("mcap: record runs past its end (truncated file?)");}
  }
  const uint8_t * p_;
  size_t n_, i_ = 0;
};

/// OMG CDR (XCDR1) reader: 4-byte encapsulation header, then primitives aligned to their own size relative to
/// the first byte after that header; strings = u32 length INCLUDING the terminating NUL; sequences = u32 count.
class Cdr
{
public:
  Cdr(const uint8_t * p, size_t n)
  : p_(p), n_(n)
  {
    if (n < 4) {throw std::runtime_error("cdr: message shorter than its encapsulation header");}
    if (p[0] != 0x00 || (p[1] != 0x01 && p[1] != 0x03)) {   // CDR_LE / PL_CDR_LE; big endian is never written on x86 / arm
      throw std::runtime_error("cdr: only little-endian CDR is supported");
    }
    i_ = 4;
  }
  template<class T> T get()
  {
    align(sizeof(T));
    need(sizeof(T));
    T v;
    std::memcpy(&v, p_ + i_, sizeof(T));
    i_ += sizeof(T);
    return v;
  }
  std::string str()
  {
    const uint32_t k = get<uint32_t>();
    need(k);
    std::string s(reinterpret_cast<const char *>(p_ + i_), k ? k - 1 : 0);
    i_ += k;
    return s;
  }
  void skip_str() {const uint32_t k = get<uint32_t>(); need(k); i_ += k;}
  size_t left() const {return n_ - i_;}
  template<class T> std::vector<T> seq()
  {
    const uint32_t k = get<uint32_t>();
    std::vector<T> v;
    if (k) {
      align(sizeof(T));
      need(size_t(k) * sizeof(T));
      v.resize(k);
      std::memcpy(v.data(), p_ + i_, size_t(k) * sizeof(T));
      i_ += size_t(k) * sizeof(T);
    }
    return v;
  }
  template<class T> void skip_seq()
  {
    const uint32_t k = get<uint32_t>();
    if (k) {align(sizeof(T)); need(size_t(k) * sizeof(T)); i_ += size_t(k) * sizeof(T);}
  }
  void skip_header() {get<int32_t>(); get<uint32_t>(); skip_str();}   // std_msgs/Header
  int64_t header_stamp_ns()
  {
    const int32_t sec = get<int32_t>();
    const uint32_t nsec = get<uint32_t>();
    skip_str();                                                         // frame_id
    return int64_t(sec) * 1000000000ll + int64_t(nsec);
  }

private:
  void align(size_t a) {const size_t rel = i_ - 4; i_ += (a - rel % a) % a;}
  void need(size_t k) const
  {
    if (i_ > n_ || k > n_ - i_) {throw std::runtime_error("cdr: field runs past the end of the message");}
  }
  const uint8_t * p_;
  size_t n_, i_ = 0;
};

/// sensor_msgs/msg/JointState: Header header; string[] name; float64[] position, velocity, effort
inline JointStateMsg decode_joint_state(const uint8_t * p, size_t n)
{
  Cdr c(p, n);
  JointStateMsg m;
  m.stamp_ns = c.header_stamp_ns();
  const uint32_t names = c.get<uint32_t>();
  for (uint32_t i = 0; i < names; ++i) {m.name.push_back(c.str());}
  m.position = c.seq<double>();
  m.velocity = c.seq<double>();
  m.effort = c.seq<double>();
  return m;
}

/// visualization_msgs/msg/MarkerArray: Marker[] markers.  Marker (ROS 2 Humble .. Jazzy): header, ns, id, type,
/// action, pose (7 f64), scale (3 f64), color (4 f32), lifetime (i32, u32), frame_locked, points[], colors[],
/// texture_resource, texture (CompressedImage: header, format, data), uv_coordinates[], text, mesh_resource,
/// mesh_file (filename, data), mesh_use_embedded_materials.
inline MarkerArrayMsg decode_marker_array(const uint8_t * p, size_t n)
{
  Cdr c(p, n);
  MarkerArrayMsg out;
  const uint32_t count = c.get<uint32_t>();
  out.markers.reserve(std::min<uint32_t>(count, 4096));
  for (uint32_t i = 0; i < count; ++i) {
    MarkerMsg m;
    c.skip_header();
    c.skip_str();                                   // ns
    m.id = c.get<int32_t>();
    m.type = c.get<int32_t>();
    m.action = c.get<int32_t>();
    m.x = c.get<double>(); m.y = c.get<double>(); m.z = c.get<double>();
    for (int q = 0; q < 4; ++q) {c.get<double>();}  // orientation
    for (int q = 0; q < 3; ++q) {m.scale[q] = c.get<double>();}
    for (int q = 0; q < 4; ++q) {c.get<float>();}   // color
    c.get<int32_t>(); c.get<uint32_t>();            // lifetime
    c.get<uint8_t>();                               // frame_locked
    {                                               // geometry_msgs/Point[] points (3 f64 each)
      const size_t k = c.get<uint32_t>();            // (counts are checked against what is left: a corrupt count
      if (k > c.left() / 24) {throw std::runtime_error("mcap: Marker.points count runs past the message");}   // must not wrap)
      for (size_t q = 0; q < k; ++q) {c.get<double>(); c.get<double>(); c.get<double>();}
    }
    {                                               // std_msgs/ColorRGBA[] colors (4 f32 each)
      const size_t k = c.get<uint32_t>();
      if (k > c.left() / 16) {throw std::runtime_error("mcap: Marker.colors count runs past the message");}
      for (size_t q = 0; q < 4 * k; ++q) {c.get<float>();}
    }
    c.skip_str();                                   // texture_resource
    c.skip_header(); c.skip_str(); c.skip_seq<uint8_t>();   // texture
    {                                               // UVCoordinate[] (2 f32 each)
      const size_t k = c.get<uint32_t>();
      if (k > c.left() / 8) {throw std::runtime_error("mcap: Marker.uv_coordinates count runs past the message");}
      for (size_t q = 0; q < 2 * k; ++q) {c.get<float>();}
    }
    c.skip_str();                                   // text
    c.skip_str();                                   // mesh_resource
    c.skip_str(); c.skip_seq<uint8_t>();            // mesh_file
    c.get<uint8_t>();                               // mesh_use_embedded_materials
    out.markers.push_back(m);
  }
  return out;
}

/// chunk decompression through the system's runtime libraries (dlopen: the image ships the .so files without headers)
inline std::vector<uint8_t> decompress_chunk(
  const std::string & method, const uint8_t * src, size_t n, uint64_t uncompressed_size)
{
  if (method.empty()) {return std::vector<uint8_t>(src, src + n);}
  // a chunk states its own uncompressed size: bound it before allocating (rosbag2's default chunk is 4 MiB; zstd / lz4
  // do not expand a block by more than a few thousand times), so that a tiny corrupt file cannot ask for gigabytes
  if (uncompressed_size > (uint64_t(1) << 30) || uncompressed_size > (uint64_t(n) + 64) * 4096) {
    throw std::runtime_error("mcap: implausible uncompressed chunk size");
  }
  std::vector<uint8_t> out(static_cast<size_t>(uncompressed_size));
  if (method == "zstd") {
    static void * lib = dlopen("libzstd.so.1", RTLD_NOW | RTLD_LOCAL);
    if (!lib) {throw std::runtime_error("mcap: chunk compression 'zstd' but libzstd.so.1 cannot be loaded");}
    using Fn = size_t (*)(void *, size_t, const void *, size_t);
    using Er = unsigned (*)(size_t);
    static Fn dec = reinterpret_cast<Fn>(dlsym(lib, "ZSTD_decompress"));
    static Er err = reinterpret_cast<Er>(dlsym(lib, "ZSTD_isError"));
    if (!dec || !err) {throw std::runtime_error("mcap: libzstd.so.1 lacks ZSTD_decompress");}
    const size_t got = dec(out.data(), out.size(), src, n);
    if (err(got) || got != out.size()) {throw std::runtime_error("mcap: zstd chunk does not decompress to its stated size");}
    return out;
  }
  if (method == "lz4") {                             // LZ4 frame format
    static void * lib = dlopen("liblz4.so.1", RTLD_NOW | RTLD_LOCAL);
    if (!lib) {throw std::runtime_error("mcap: chunk compression 'lz4' but liblz4.so.1 cannot be loaded");}
    using Create = size_t (*)(void **, unsigned);
    using Free = size_t (*)(void *);
    using Dec = size_t (*)(void *, void *, size_t *, const void *, size_t *, const void *);
    using Er = unsigned (*)(size_t);
    static Create create = reinterpret_cast<Create>(dlsym(lib, "LZ4F_createDecompressionContext"));
    static Free release = reinterpret_cast<Free>(dlsym(lib, "LZ4F_freeDecompressionContext"));
    static Dec dec = reinterpret_cast<Dec>(dlsym(lib, "LZ4F_decompress"));
    static Er err = reinterpret_cast<Er>(dlsym(lib, "LZ4F_isError"));
    if (!create || !release || !dec || !err) {throw std::runtime_error("mcap: liblz4.so.1 lacks the LZ4F frame API");}
    void * ctx = nullptr;
    if (err(create(&ctx, 100))) {throw std::runtime_error("mcap: LZ4F context");}   // LZ4F_VERSION = 100
    size_t done_in = 0, done_out = 0;
    bool bad = false;
    while (done_in < n) {
      size_t dn = out.size() - done_out, sn = n - done_in;
      const size_t rc = dec(ctx, out.data() + done_out, &dn, src + done_in, &sn, nullptr);
      if (err(rc) || (dn == 0 && sn == 0)) {bad = true; break;}
      done_in += sn; done_out += dn;
      if (rc == 0) {break;}
    }
    release(ctx);
    if (bad || done_out != out.size()) {throw std::runtime_error("mcap: lz4 chunk does not decompress to its stated size");}
    return out;
  }
  throw std::runtime_error("mcap: unknown chunk compression '" + method + "'");
}

/// One parsed bag: the joint-state and marker-array messages of the requested topics, each list in log-time
/// order, and the merged event order the node's single-threaded executor would see (slam.cpp:680-686).
class McapReader
{
public:
  struct Event {uint64_t log_time; int32_t kind; int32_t index;};   ///< kind 0: joint_states[index], 1: detections[index]

  std::vector<JointStateMsg> joint_states;
  std::vector<MarkerArrayMsg> detections;
  std::vector<Event> events;
  std::string profile, library;
  uint64_t messages_total = 0, chunks = 0;
  std::map<std::string, uint64_t> topic_counts;         ///< every topic of the file -> messages seen

  McapReader(
    const std::string & path, const std::string & joint_topic = "/blue/joint_states",
    const std::string & detections_topic = "/green/detected_obstacles")
  : joint_topic_(joint_topic), det_topic_(detections_topic)
  {
    std::vector<uint8_t> file = slurp(path);
    static const uint8_t magic[8] = {0x89, 'M', 'C', 'A', 'P', 0x30, '\r', '\n'};
    if (file.size() < 16 || std::memcmp(file.data(), magic, 8) != 0) {
      throw std::runtime_error("mcap: '" + path + "' does not start with the MCAP magic");
    }
    walk(file.data() + 8, file.size() - 8, false);
    auto by_time = [](const auto & a, const auto & b) {return a.log_time < b.log_time;};
    std::stable_sort(joint_states.begin(), joint_states.end(), by_time);
    std::stable_sort(detections.begin(), detections.end(), by_time);
    for (size_t i = 0; i < joint_states.size(); ++i) {events.push_back({joint_states[i].log_time, 0, int32_t(i)});}
    for (size_t i = 0; i < detections.size(); ++i) {events.push_back({detections[i].log_time, 1, int32_t(i)});}
    std::stable_sort(events.begin(), events.end(), by_time);   // ties: joint state before detections
  }

private:
  static std::vector<uint8_t> slurp(const std::string & path)
  {
    FILE * f = std::fopen(path.c_str(), "rb");
    if (!f) {throw std::runtime_error("mcap: cannot open '" + path + "'");}
    std::vector<uint8_t> buf;
    uint8_t tmp[1 << 16];
    size_t k;
    while ((k = std::fread(tmp, 1, sizeof(tmp), f)) > 0) {buf.insert(buf.end(), tmp, tmp + k);}
    std::fclose(f);
    return buf;
  }

  /// records of the data section (or of one chunk); stops at DataEnd / Footer
  void walk(const uint8_t * p, size_t n, bool in_chunk)
  {
    Cursor c(p, n);
    while (c.left() >= 9) {
      const uint8_t op = c.get<uint8_t>();
      const uint64_t len = c.get<uint64_t>();
      if (len > c.left()) {throw std::runtime_error("mcap: record runs past its end (truncated file?)");}
      Cursor r(c.here(), static_cast<size_t>(len));
      c.skip(static_cast<size_t>(len));
      switch (op) {
        case 0x01: profile = r.str32(); library = r.str32(); break;             // Header
        case 0x04: {                                                            // Channel
            const uint16_t id = r.get<uint16_t>();
            r.get<uint16_t>();                                                  // schema id
            const std::string topic = r.str32();
            const std::string encoding = r.str32();
            channels_[id] = topic;
            if ((topic == joint_topic_ || topic == det_topic_) && encoding != "cdr") {
              throw std::runtime_error("mcap: topic " + topic + " is encoded as '" + encoding + "', expected cdr");
            }
            break;
          }
        case 0x05: {                                                            // Message
            const uint16_t ch = r.get<uint16_t>();
            r.get<uint32_t>();                                                  // sequence
            const uint64_t log_time = r.get<uint64_t>();
            r.get<uint64_t>();                                                  // publish time
            ++messages_total;
            auto it = channels_.find(ch);
            if (it == channels_.end()) {throw std::runtime_error("mcap: message on an undeclared channel");}
            ++topic_counts[it->second];
            if (it->second == joint_topic_) {
              joint_states.push_back(decode_joint_state(r.here(), r.left()));
              joint_states.back().log_time = log_time;
            } else if (it->second == det_topic_) {
              detections.push_back(decode_marker_array(r.here(), r.left()));
              detections.back().log_time = log_time;
            }
            break;
          }
        case 0x06: {                                                            // Chunk
            if (in_chunk) {throw std::runtime_error("mcap: chunk inside a chunk");}
            r.get<uint64_t>(); r.get<uint64_t>();                               // message start / end time
            const uint64_t usize = r.get<uint64_t>();
            r.get<uint32_t>();                                                  // crc of the uncompressed records
            const std::string method = r.str32();
            const uint64_t csize = r.get<uint64_t>();
            if (csize > r.left()) {throw std::runtime_error("mcap: chunk payload runs past its record");}
            const std::vector<uint8_t> records = decompress_chunk(method, r.here(), static_cast<size_t>(csize), usize);
            ++chunks;
            walk(records.data(), records.size(), true);
            break;
          }
        case 0x0F: case 0x02: return;                                           // DataEnd / Footer
        default: break;   // Schema, MessageIndex, ChunkIndex, Attachment, Statistics, Metadata, SummaryOffset: not needed
      }
    }
  }

  std::string joint_topic_, det_topic_;
  std::map<uint16_t, std::string> channels_;
};

}  // namespace rosbag
#endif
