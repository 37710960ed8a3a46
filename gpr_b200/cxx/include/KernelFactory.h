// String <-> kernel tree, interface-compatible with the reference's include/KernelFactory.h:
//   KernelFactory<T>::GetKernel(std::string&)   kernel_string of the JSON config and of
//                                               "-ParameterFile.txt" (reference :83-178), e.g.
//                                               "GaussianKernel(35, 30,)" or nested Sum/Product forms
//   KernelFactory<T>::Load(name, string_params) registry lookup (:67-70)
//   KernelFactory<T>::RegisterKernels / AddType (:61-80)
// Implementation: a name -> loader table and a recursive-descent parse of the ToString() grammar
//   kernel := Name '(' (kernel ',' kernel | param ',' ...) ')'
// As in the reference, GetKernel consumes its argument: on return the string holds the unparsed rest.
#ifndef GPRB200_CXX_KERNEL_FACTORY_H
#define GPRB200_CXX_KERNEL_FACTORY_H

#include <map>
#include <memory>
#include <string>

#include "Kernel.h"

namespace gpr {

template <typename TKernel, class TScalarType>
std::shared_ptr<Kernel<TScalarType> > CreateKernelFromStrings(
    const typename Kernel<TScalarType>::StringParameterVectorType& parameters) {
    return TKernel::Load(parameters);
}

template <class TScalarType>
class KernelFactory {
public:
    typedef Kernel<TScalarType> KernelType;
    typedef std::shared_ptr<KernelType> KernelTypePointer;
    typedef KernelTypePointer (*ComponentFactoryFuncPtr)(const typename KernelType::StringParameterVectorType&);
    typedef std::map<const std::string, ComponentFactoryFuncPtr> MapType;

    static std::shared_ptr<MapType> m_Map;

    template <typename TKernel>
    static void AddType(const std::string& componentName) {
        Registry()[componentName] = &CreateKernelFromStrings<TKernel, TScalarType>;
    }

    static KernelTypePointer Load(const std::string& kernel_string,
                                  const typename KernelType::StringParameterVectorType& parameters) {
        typename MapType::iterator it = Registry().find(kernel_string);
        if (it == Registry().end()) throw std::string("KernelFactory::GetKernel: failed to load kernel.");
        return (*it->second)(parameters);
    }

    static void RegisterKernels() {
        if (!Registry().empty()) return;
        AddType<WhiteKernel<TScalarType> >("WhiteKernel");
        AddType<GaussianKernel<TScalarType> >("GaussianKernel");
        AddType<GaussianExpKernel<TScalarType> >("GaussianExpKernel");
        AddType<PeriodicKernel<TScalarType> >("PeriodicKernel");
        AddType<RationalQuadraticKernel<TScalarType> >("RationalQuadraticKernel");
        AddType<SumKernel<TScalarType> >("SumKernel");
        AddType<ProductKernel<TScalarType> >("ProductKernel");
    }

    static KernelTypePointer GetKernel(std::string& kernel_string) {
        RegisterKernels();
        size_t pos = 0;
        KernelTypePointer k = Parse(kernel_string, pos);
        kernel_string = kernel_string.substr(pos);
        return k;
    }

private:
    static MapType& Registry() {
        if (!m_Map) m_Map.reset(new MapType());
        return *m_Map;
    }

    static void SkipSpace(const std::string& s, size_t& pos) {
        while (pos < s.size() && (s[pos] == ' ' || s[pos] == '\t')) ++pos;
    }

    // parses one kernel starting at pos; leaves pos just behind its closing parenthesis
    static KernelTypePointer Parse(const std::string& s, size_t& pos) {
        SkipSpace(s, pos);
        const size_t open = s.find('(', pos);
        if (open == std::string::npos) throw std::string("KernelFactory::GetKernel: failed to tokanize kernel name string");
        const std::string name = s.substr(pos, open - pos);
        pos = open + 1;
        const bool is_sum = (name == "SumKernel"), is_product = (name == "ProductKernel");
        if (is_sum || is_product) {
            KernelTypePointer k1 = Parse(s, pos);
            SkipSpace(s, pos);
            if (pos >= s.size() || s[pos] != ',')
                throw std::string(is_sum ? "KernelFactory::GetKernel: failed to tokanize  sum kernel name string"
                                         : "KernelFactory::GetKernel: failed to tokanize  product kernel name string");
            ++pos;
            KernelTypePointer k2 = Parse(s, pos);
            SkipSpace(s, pos);
            if (pos < s.size() && s[pos] == ')') ++pos;
            if (is_sum) return KernelTypePointer(new SumKernel<TScalarType>(k1, k2));
            return KernelTypePointer(new ProductKernel<TScalarType>(k1, k2));
        }
        // leaf: comma-terminated parameters up to the closing parenthesis ("35, 30,)" or "35,30)")
        typename KernelType::StringParameterVectorType parameters;
        std::string token;
        for (; pos < s.size(); ++pos) {
            const char c = s[pos];
            if (c == ',' || c == ')') {
                if (token.find_first_not_of(" \t") != std::string::npos) parameters.push_back(token);
                token.clear();
                if (c == ')') {
                    ++pos;
                    break;
                }
            } else {
                token += c;
            }
        }
        if (Registry().find(name) == Registry().end()) throw std::string("KernelFactory::GetKernel: failed to load kernel.");
        return Load(name, parameters);
    }
};

template <class TScalarType>
std::shared_ptr<typename KernelFactory<TScalarType>::MapType> KernelFactory<TScalarType>::m_Map;

}  // namespace gpr

#endif
