// Event<T> / EventHandler<T> — the reference's callback list (core/src/utils/event.h:6-49): handlers are plain
// function pointers with a context; emit() walks a snapshot of the list so a handler may unbind itself.
#pragma once
#include <algorithm>
#include <cstdio>
#include <vector>

template <class T>
struct EventHandler {
    EventHandler() {}
    EventHandler(void (*h)(T, void*), void* c) : handler(h), ctx(c) {}
    void (*handler)(T, void*) = nullptr;
    void* ctx = nullptr;
};

template <class T>
class Event {
public:
    void emit(T value) {
        const std::vector<EventHandler<T>*> snapshot = handlers;
        for (auto* h : snapshot) { h->handler(value, h->ctx); }
    }

    void bindHandler(EventHandler<T>* h) { handlers.push_back(h); }

    void unbindHandler(EventHandler<T>* h) {
        auto it = std::find(handlers.begin(), handlers.end(), h);
        if (it == handlers.end()) { fprintf(stderr, "Tried to remove a non-existent event handler\n"); return; }
        handlers.erase(std::remove(handlers.begin(), handlers.end(), h), handlers.end());
    }

    std::vector<EventHandler<T>*> handlers;
};
