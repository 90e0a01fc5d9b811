// dsp::chain<T> — an ordered list of Processor<T,T> blocks, each of which can be switched in or out while the
// graph runs (reference: core/src/dsp/chain.h:8-194).  Same public surface: init / setInput / addBlock / removeBlock /
// enableBlock / disableBlock / setBlockEnabled / enableAllBlocks / disableAllBlocks / start / stop and the public
// `out` pointer, which always names the stream a consumer must read (the chain's input when nothing is enabled).
// Implementation: after every change the enabled blocks are re-linked in list order (block k reads the output of
// the previous enabled block), and `onOutputChange(out)` is invoked when the tail stream changed.
#pragma once
#include <algorithm>
#include <map>
#include <stdexcept>
#include <vector>

#include "processor.h"

namespace dsp {
    template <class T>
    class chain {
    public:
        chain() {}
        explicit chain(stream<T>* in) { init(in); }

        void init(stream<T>* in) {
            _in = in;
            out = _in;
        }

        template <typename Func>
        void setInput(stream<T>* in, Func onOutputChange) {
            _in = in;
            relink(onOutputChange);
        }

        void addBlock(Processor<T, T>* block, bool enabled) {
            if (blockExists(block)) { throw std::runtime_error("[chain] Tried to add a block that is already part of the chain"); }
            links.push_back(block);
            states[block] = false;
            if (enabled) { enableBlock(block, [](stream<T>*) {}); }
        }

        template <typename Func>
        void removeBlock(Processor<T, T>* block, Func onOutputChange) {
            if (!blockExists(block)) { throw std::runtime_error("[chain] Tried to remove a block that is not part of the chain"); }
            disableBlock(block, onOutputChange);
            states.erase(block);
            links.erase(std::find(links.begin(), links.end(), block));
        }

        template <typename Func>
        void enableBlock(Processor<T, T>* block, Func onOutputChange) {
            if (!blockExists(block)) { throw std::runtime_error("[chain] Tried to enable a block that isn't part of the chain"); }
            if (states[block]) { return; }
            states[block] = true;
            relink(onOutputChange);
            if (running) { block->start(); }
        }

        template <typename Func>
        void disableBlock(Processor<T, T>* block, Func onOutputChange) {
            if (!blockExists(block)) { throw std::runtime_error("[chain] Tried to disable a block that isn't part of the chain"); }
            if (!states[block]) { return; }
            block->stop();
            states[block] = false;
            relink(onOutputChange);
        }

        template <typename Func>
        void setBlockEnabled(Processor<T, T>* block, bool enabled, Func onOutputChange) {
            if (enabled) { enableBlock(block, onOutputChange); }
            else { disableBlock(block, onOutputChange); }
        }

        template <typename Func>
        void enableAllBlocks(Func onOutputChange) {
            for (auto* b : links) { enableBlock(b, onOutputChange); }
        }

        template <typename Func>
        void disableAllBlocks(Func onOutputChange) {
            for (auto* b : links) { disableBlock(b, onOutputChange); }
        }

        void start() {
            if (running) { return; }
            for (auto* b : links) { if (states[b]) { b->start(); } }
            running = true;
        }

        void stop() {
            if (!running) { return; }
            for (auto* b : links) { if (states[b]) { b->stop(); } }
            running = false;
        }

        bool blockExists(Processor<T, T>* block) { return states.find(block) != states.end(); }

        stream<T>* out = nullptr;

    private:
        // re-plumb: each enabled block reads the previous enabled block's output (or the chain input)
        template <typename Func>
        void relink(Func onOutputChange) {
            stream<T>* cur = _in;
            for (auto* b : links) {
                if (!states[b]) { continue; }
                b->setInput(cur);
                cur = &b->out;
            }
            if (cur != out) {
                out = cur;
                onOutputChange(out);
            }
        }

        stream<T>* _in = nullptr;
        std::vector<Processor<T, T>*> links;
        std::map<Processor<T, T>*, bool> states;
        bool running = false;
    };
}
