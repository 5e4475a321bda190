// LocalFrame.h — an image pair; the adjuster reads getImage2() of every frame (mirror of src/frame/LocalFrame.h:19-83).
#ifndef B200BA_LOCAL_FRAME_H
#define B200BA_LOCAL_FRAME_H
#include <memory>

#include "../component/Image.h"

class LocalFrame {
public:
    using Ptr = std::shared_ptr<LocalFrame>;

private:
    Image::Ptr image1_ptr_, image2_ptr_;

public:
    [[nodiscard]] LocalFrame(const std::shared_ptr<Image>& image1_ptr, const std::shared_ptr<Image>& image2_ptr)
        : image1_ptr_(image1_ptr), image2_ptr_(image2_ptr) {}
    auto getImage1() { return image1_ptr_; }
    auto getImage2() const { return image2_ptr_; }
};
#endif
