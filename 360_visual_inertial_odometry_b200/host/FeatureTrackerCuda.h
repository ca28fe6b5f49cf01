// FeatureTrackerCuda.h -- drop-in for the two hot methods of vio_360::FeatureTracker
// (reference: src/processing/FeatureTracker.h:54-83, src/processing/FeatureTracker.cpp).
//
//   std::vector<cv::Point2f> DetectNewFeatures(const cv::Mat& image, const cv::Mat& mask = cv::Mat());
//   int TrackOpticalFlow(const cv::Mat& prev, const cv::Mat& cur, const std::vector<cv::Point2f>& prev_pts,
//                        std::vector<cv::Point2f>& cur_pts, std::vector<uchar>& status);
//
// Same names, argument meaning and return values; the arithmetic runs in libvio360_cuda.so
// through the C ABI (include/vio360.h).  There is no CPU fallback: construction throws
// std::runtime_error when no B200-class device is usable, and every call throws on a C-ABI
// error (the reference itself has no error channel; OpenCV would throw cv::Exception).
// INTEGRATION.md shows the three-line change in FeatureTracker.cpp that routes the
// reference through this class.
#pragma once

#include <cstdint>
#include <memory>
#include <random>
#include <string>
#include <vector>

#include "../../include/vio360.h"
#include "cv_compat.h"

namespace vio_360 {

class FeatureTrackerCuda {
public:
    // Mirrors FeatureTracker::FeatureTracker (FeatureTracker.cpp:24-59).  `camera` is anything with
    // GetWidth()/GetHeight() (the reference's std::shared_ptr<Camera>).
    template <class CameraPtr>
    explicit FeatureTrackerCuda(CameraPtr camera, int max_features = 500, float min_distance = 30.0f,
                                float quality_level = 0.01f, int boundary_margin = 20, int device = 0)
        : FeatureTrackerCuda(camera->GetWidth(), camera->GetHeight(), max_features, min_distance, quality_level,
                             boundary_margin, device) {}
    FeatureTrackerCuda(int width, int height, int max_features = 500, float min_distance = 30.0f,
                       float quality_level = 0.01f, int boundary_margin = 20, int device = 0);
    ~FeatureTrackerCuda();
    FeatureTrackerCuda(const FeatureTrackerCuda&) = delete;
    FeatureTrackerCuda& operator=(const FeatureTrackerCuda&) = delete;

    // FeatureTracker.cpp:208-226
    std::vector<cv::Point2f> DetectNewFeatures(const cv::Mat& image, const cv::Mat& mask = cv::Mat());
    // DetectNewFeatures(image, CreateFeatureMask(features)) (FeatureTracker.cpp:177-178, 386-402) without
    // rasterising and uploading the mask: pass the pixel coordinates of the valid existing features; the
    // discs cv::circle(mask, pt, int(min_distance), 0, -1) are applied on the device.
    std::vector<cv::Point2f> DetectNewFeatures(const cv::Mat& image, const std::vector<cv::Point2f>& occupied);
    // FeatureTracker.cpp:228-251
    int TrackOpticalFlow(const cv::Mat& prev_image, const cv::Mat& curr_image,
                         const std::vector<cv::Point2f>& prev_points, std::vector<cv::Point2f>& curr_points,
                         std::vector<uchar>& status);
    // FeatureTracker.cpp:253-328.  The 3-point samples are drawn here exactly as the reference draws them
    // (std::mt19937 over std::uniform_int_distribution, repeats redrawn, :273-289); hypotheses and inlier
    // counting run on the device.  SetRansacSeed replaces the std::random_device seed (tests).
    int RejectOutliersRotationRANSAC(const std::vector<cv::Point2f>& prev_points,
                                     const std::vector<cv::Point2f>& curr_points, std::vector<uchar>& inlier_mask);
    void SetRansacSeed(uint32_t seed) { m_ransac_seed = seed; m_ransac_seeded = true; }
    void SetRansacParams(int max_iterations, float threshold_deg) { m_ransac_max_iterations = max_iterations; m_ransac_threshold = threshold_deg; }
    // FeatureTracker::RemoveClusteredFeatures (FeatureTracker.cpp:404-497) over the features' pixel coordinates:
    // keep[i] == 0 for features of clustered grid cells (the caller erases them from the Frame in order).
    // Grid and ratio default to the reference's (FeatureTracker.cpp:39-40, default_config.yaml:54).
    std::vector<uchar> RemoveClusteredFeatures(const std::vector<cv::Point2f>& points, float clustered_std_ratio = 0.25f,
                                               int grid_cols = 20, int grid_rows = 10);
    // Frame::AssignFeaturesToGrid + Frame::LimitFeaturesPerGrid (Frame.cpp:108-202): keep[i] == 1 for the features that
    // survive the per-cell cap (highest track counts first, ties in feature order).
    std::vector<uchar> LimitFeaturesPerGrid(const std::vector<cv::Point2f>& points, const std::vector<int>& track_counts,
                                            int max_features_per_grid = 4, int grid_cols = 20, int grid_rows = 10);
    // FeatureTracker::TrackFeatures (FeatureTracker.cpp:61-206) for the consecutive frames of one sequence as ONE
    // device-resident call per frame (vio360_fe_track_frame_batch): tracking, post filter, rotation RANSAC, clustered
    // cell removal, grid cap, re-detection under the occupancy mask, second grid cap.  The feature list lives on
    // the device between the frames; what comes back is what Frame::GetFeatures() would hold: ids (carried from the
    // previous frame or fresh from m_next_feature_id), pixel coordinates, track counts, in the reference's order.
    // The shared_ptr<Feature> / observation bookkeeping (:139-163) stays with the caller.  The first call (or the
    // first after ResetSequence) is the "no previous frame" branch (:68-92).
    struct FrameFeatures {
        std::vector<int> ids; std::vector<cv::Point2f> points; std::vector<int> track_counts;
        int num_tracked = 0, num_detected = 0;
    };
    void TrackFeatures(const cv::Mat& image, FrameFeatures& out);
    void ResetSequence();
    // Frame::SetGridParameters (Estimator.cpp:572-576) and the ConfigUtils switches RemoveClusteredFeatures reads
    // (FeatureTracker.cpp:407-423); call before the first TrackFeatures.  Defaults: default_config.yaml.
    void SetGridParameters(int grid_cols, int grid_rows, int max_features_per_grid) { m_grid_cols = grid_cols; m_grid_rows = grid_rows; m_grid_cap = max_features_per_grid; ResetSequence(); }
    void SetClusterRemoval(bool enabled, float std_ratio) { m_cluster = enabled; m_cluster_ratio = std_ratio; ResetSequence(); }

    // The post filter of TrackFeatures (FeatureTracker.cpp:118-126) computed on the device
    // during the last TrackOpticalFlow call: status && !polar && !near-boundary.
    const std::vector<uchar>& LastKeepMask() const { return m_keep; }

    // Frame identity for the device pyramid cache: when the caller tags images (e.g. with the
    // Frame id) the pyramid of `prev` is re-used instead of being rebuilt every call as
    // cv::calcOpticalFlowPyrLK does.  0 = no caching (exact reference behaviour, slower).
    void SetFrameKeys(uint64_t prev_key, uint64_t curr_key) { m_prev_key = prev_key; m_curr_key = curr_key; }

    void SetMaxFeatures(int max_features) { m_max_features = max_features; }
    void SetMinDistance(float min_distance) { m_min_distance = min_distance; }
    int GetMaxFeatures() const { return m_max_features; }
    void GetTrackingStats(int& num_tracked, int& num_detected) const { num_tracked = m_num_tracked; num_detected = m_num_detected; }

    // Host frame buffers.  The reference's images are pageable cv::Mat buffers (cv::imread / the caller's ring);
    // copying from pageable memory runs at a fraction of the PCIe rate.  RegisterFrameBuffer page-locks the Mat's
    // buffer in place (vio360_host_register) -- call it once per buffer of a camera ring.  With
    // SetAutoRegisterFrames(true) every image handed to TrackFeatures / TrackOpticalFlow / DetectNewFeatures is
    // registered the first time its buffer is seen; at most `max_buffers` stay registered (least recently used first
    // out), so a caller that allocates a fresh Mat per frame pays one registration per frame instead of a staged copy.
    void RegisterFrameBuffer(const cv::Mat& image);
    void UnregisterFrameBuffer(const cv::Mat& image);
    void SetAutoRegisterFrames(bool on, size_t max_buffers = 8) { m_auto_register = on; m_max_registered = max_buffers; }

private:
    void check(int rc, const char* what) const;
    vio360_ctx* m_ctx = nullptr;
    int m_width, m_height, m_max_features;
    float m_min_distance, m_quality_level;
    int m_boundary_margin;
    int m_point_capacity = 0;
    uint64_t m_prev_key = 0, m_curr_key = 0;
    int m_num_tracked = 0, m_num_detected = 0;
    int m_ransac_max_iterations = 1000;   // FeatureTracker.cpp:36
    float m_ransac_threshold = 2.0f;      // degrees, FeatureTracker.cpp:37
    uint32_t m_ransac_seed = 0;
    bool m_ransac_seeded = false;
    std::vector<uchar> m_keep;
    vio360_seq* m_seq = nullptr;
    int m_grid_cols = 20, m_grid_rows = 10, m_grid_cap = 10;   // default_config.yaml:14-16
    bool m_cluster = true; float m_cluster_ratio = 0.25f;      // default_config.yaml:52-54
    std::mt19937* m_seq_rng = nullptr;
    void ensure_capacity(int n_points);
    void touch_frame(const cv::Mat& image);                 // auto registration + LRU bookkeeping
    bool m_auto_register = false;
    size_t m_max_registered = 8;
    std::vector<std::pair<const void*, size_t>> m_registered;   // most recently used last
};

}  // namespace vio_360
