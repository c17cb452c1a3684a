// channels.hpp -- the two hand-off structures of the host pipelines (text and dazcon): a bounded FIFO between stages
// and an outbox that releases finished batches in input order.  They stand where the reference has its
// BoundedBuffer (reference src/cpp/BoundedBuffer.hpp:18-50), with batches of many targets as the unit.
#pragma once
#include <condition_variable>
#include <cstdint>
#include <deque>
#include <map>
#include <memory>
#include <mutex>

namespace dgh {

template <class T>
class Channel {  // bounded FIFO; close() wakes everybody
public:
    explicit Channel(size_t cap) : cap_(cap) {}
    bool push(T v) {
        std::unique_lock<std::mutex> lk(mu_);
        cv_space_.wait(lk, [&] { return q_.size() < cap_ || closed_; });
        if (closed_) return false;
        q_.push_back(std::move(v));
        cv_item_.notify_one();
        return true;
    }
    bool pop(T& v) {
        std::unique_lock<std::mutex> lk(mu_);
        cv_item_.wait(lk, [&] { return !q_.empty() || closed_; });
        if (q_.empty()) return false;
        v = std::move(q_.front());
        q_.pop_front();
        cv_space_.notify_one();
        return true;
    }
    bool try_pop(T& v) {
        std::lock_guard<std::mutex> lk(mu_);
        if (q_.empty()) return false;
        v = std::move(q_.front());
        q_.pop_front();
        cv_space_.notify_one();
        return true;
    }
    void close() {
        std::lock_guard<std::mutex> lk(mu_);
        closed_ = true;
        cv_item_.notify_all();
        cv_space_.notify_all();
    }

private:
    std::mutex mu_;
    std::condition_variable cv_item_, cv_space_;
    std::deque<T> q_;
    size_t cap_;
    bool closed_ = false;
};

// Finished items come out of the workers in any order; the consumer takes them by sequence number.
template <class B>
class OrderedOutboxT {
public:
    void put(std::unique_ptr<B> b) {
        std::lock_guard<std::mutex> lk(mu_);
        const uint64_t s = b->seq;
        ready_[s] = std::move(b);
        cv_.notify_all();
    }
    // nullptr when the stream is finished (or failed)
    std::unique_ptr<B> take(uint64_t seq) {
        std::unique_lock<std::mutex> lk(mu_);
        cv_.wait(lk, [&] { return ready_.count(seq) || done_; });
        auto it = ready_.find(seq);
        if (it == ready_.end()) return nullptr;
        auto b = std::move(it->second);
        ready_.erase(it);
        return b;
    }
    void finish() {
        std::lock_guard<std::mutex> lk(mu_);
        done_ = true;
        cv_.notify_all();
    }

private:
    std::mutex mu_;
    std::condition_variable cv_;
    std::map<uint64_t, std::unique_ptr<B>> ready_;
    bool done_ = false;
};

}  // namespace dgh
